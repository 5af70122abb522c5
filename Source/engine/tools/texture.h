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

// Set_Pixel (texture.h:27-38) for host code that draws into the bitmap itself (overlays, test patterns): {r, g, b} bytes to the
// 32-bit B,G,R,0 pixel at column x of row y (row 0 = bottom). The renderer's own pixels are packed the same way on the device.
inline void Set_Pixel(const Vec3<uint8>& color, Texture& texture, int32 x, int32 y)
{
    if (texture.file_type != TextureFileType::BMP) return;
    if (x < 0 || y < 0 || (uint32)x >= texture.bmb.width || (uint32)y >= texture.bmb.height) return;
    ((uint32*)texture.bmb.buffer_memory)[(size_t)y * texture.bmb.width + (size_t)x] = ((uint32)color.x << 16) | ((uint32)color.y << 8) | (uint32)color.z;
}

// Write_To_File (texture.cpp:66-115, 147-158): byte-identical BMP into "<file_name>_<id>.bmp", first id not yet on disk.
inline bool Write_To_File(Texture& tex, const char* file_name)
{
    if (tex.file_type != TextureFileType::BMP) return false;
    return atray_write_bmp((const uint32_t*)tex.bmb.buffer_memory, tex.bmb.width, tex.bmb.height, file_name, nullptr, 0) == ATRAY_OK;
}
