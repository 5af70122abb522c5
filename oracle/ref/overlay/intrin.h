// Linux stand-in for MSVC <intrin.h>, used ONLY to compile the reference's hot path into
// oracle/_ref (test infrastructure). Defines the few MSVC/SVML intrinsics PL_math.h uses.
#pragma once
#include <x86intrin.h>
#include <math.h>
// SVML _mm_invsqrt_ps (PL_math.h:390) is not available outside MSVC/ICC. Project definition
// (SURVEY.md §0-7): 1.0f / sqrtf(x), two IEEE round-to-nearest operations.
static inline __m128 _mm_invsqrt_ps(__m128 v) { return _mm_set_ss(1.0f / sqrtf(_mm_cvtss_f32(v))); }
// The reference casts int32* to long* (32-bit on Win64, 64-bit here): operate on 32 bits.
static inline int _InterlockedExchangeAdd(volatile long* p, int v) { return __sync_fetch_and_add((volatile int*)p, v); }
static inline int _InterlockedIncrement(volatile long* p) { return __sync_add_and_fetch((volatile int*)p, 1); }
static inline int _InterlockedDecrement(volatile long* p) { return __sync_sub_and_fetch((volatile int*)p, 1); }
static inline int _InterlockedExchange(volatile long* p, int v) { return __atomic_exchange_n((volatile int*)p, v, __ATOMIC_SEQ_CST); }
static inline long long _InterlockedExchangeAdd64(volatile long long* p, long long v) { return __sync_fetch_and_add(p, v); }
static inline long long _InterlockedIncrement64(volatile long long* p) { return __sync_add_and_fetch(p, 1); }
static inline long long _InterlockedDecrement64(volatile long long* p) { return __sync_sub_and_fetch(p, 1); }
static inline long long _InterlockedExchange64(volatile long long* p, long long v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
#ifndef __debugbreak
#define __debugbreak() __builtin_trap()
#endif
