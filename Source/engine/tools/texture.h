// Texture / BitmapBuffer (engine/tools/texture.h:5-42): 32-bit BGRX, bottom-up. Setup_Texture allocates PINNED host
// memory so the renderer's device→host copy of the finished bitmap is a direct async DMA.
#pragma once
#include <cstdio>
#include <cstring>
#include "PL/PL_math.h"
#include "atray_b200.h"
enum struct TextureFileType { BMP };
struct BitmapBuffer
{
    void* buffer_memory = nullptr;
    uint32 bytes_per_pixel = 4, width = 0, height = 0;
    inline uint32 size() { return bytes_per_pixel * width * height; }
};
struct Texture { TextureFileType file_type = TextureFileType::BMP; BitmapBuffer bmb; };

inline bool Setup_Texture(Texture& tex, TextureFileType file_type, uint32 width, uint32 height)  // texture.cpp:133-146
{
    tex.file_type = file_type;
    tex.bmb.bytes_per_pixel = 4; tex.bmb.width = width; tex.bmb.height = height;
    if (atray_host_alloc_pinned((size_t)width * height * 4, &tex.bmb.buffer_memory) != ATRAY_OK) return false;
    std::memset(tex.bmb.buffer_memory, 0, (size_t)width * height * 4);
    return true;
}
inline void Free_Texture(Texture& tex) { if (tex.bmb.buffer_memory) atray_host_free_pinned(tex.bmb.buffer_memory); tex.bmb.buffer_memory = nullptr; }

// Write_To_File (texture.cpp:66-115, 147-158): byte-identical BMP into "<file_name>_<id>.bmp", first id not yet on disk.
inline bool Write_To_File(Texture& tex, const char* file_name)
{
    if (tex.file_type != TextureFileType::BMP) return false;
    return atray_write_bmp((const uint32_t*)tex.bmb.buffer_memory, tex.bmb.width, tex.bmb.height, file_name, nullptr, 0) == ATRAY_OK;
}
