// Minimal base types for the host layer (fresh; names follow the reference's PL/PL_base_defs.h:52-66 and
// PL/PL_math.h:83-129 so code written against the reference's renderer API compiles unchanged).
#pragma once
#include <cstdint>
#include <cstddef>

typedef int8_t int8; typedef int16_t int16; typedef int32_t int32; typedef long long int64;
typedef uint8_t uint8; typedef uint16_t uint16; typedef uint32_t uint32; typedef unsigned long long uint64;
typedef bool b8; typedef int b32; typedef float f32; typedef double f64;
#ifndef TRUE
#define TRUE 1
#endif
#ifndef FALSE
#define FALSE 0
#endif

template <typename t> struct Vec2 { t x, y; };
template <typename t> struct Vec3 { t x, y, z; };
template <typename t> struct Vec4 { Vec2<t> left_bottom, right_top; };
typedef Vec2<int32> vec2i;
typedef Vec3<f32> vec3f;
typedef Vec3<uint8> vec3b;
typedef Vec4<int32> Tile;  // PL_math.h:185

// Growable array with the reference's DBuffer surface (pl_utils.h:27-98): add / add_nocpy / length / front / [].
template <typename t> struct DBuffer
{
    int32 length = 0, capacity = 0;
    t* front = nullptr;
    t* add(t v) { return add_nocpy(v); }
    t* add_nocpy(t& v)
    {
        if (length == capacity)
        {
            int32 nc = capacity ? capacity * 2 : 4;
            t* n = new t[nc];
            for (int32 i = 0; i < length; ++i) n[i] = front[i];
            delete[] front;
            front = n;
            capacity = nc;
        }
        front[length] = v;
        return &front[length++];
    }
    void clear_buffer() { delete[] front; front = nullptr; length = capacity = 0; }
    t& operator[](int32 i) { return front[i]; }
};
// Fixed array (pl_utils.h:102-146)
template <typename t, typename size_type = int32> struct FDBuffer
{
    size_type size = 0;
    t* front = nullptr;
    t* allocate(size_type n) { size = n; front = new t[n](); return front; }
    void clear() { delete[] front; front = nullptr; size = 0; }
    t& operator[](size_type i) { return front[i]; }
};
