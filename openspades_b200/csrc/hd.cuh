// Host/device wrappers for the handful of intrinsics the terrain core uses, so that terrain_core.cuh compiles three
// ways: nvcc device pass (the product), nvcc host pass, and plain g++ (tests/native/terrain_emul.cpp, which runs the
// very same per-lane code sequentially on the CPU against the oracle -- this container has no GPU, so control-flow
// and arithmetic mistakes are caught here before GPU minutes are spent).
//
// Every float operation is a single IEEE rounding (the reference is x86-64 SSE scalar code without FMA contraction);
// the product is built with --fmad=false and the host harness with -ffp-contract=off.
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define OSB_HDI __host__ __device__ __forceinline__
#define OSB_HDM __host__ __device__ __forceinline__ // member functions
#else
#include <cmath>
#define OSB_HDI static inline
#define OSB_HDM inline
#ifndef __restrict__
#define __restrict__
#endif
struct float2 {
	float x, y;
};
struct uint2 {
	uint32_t x, y;
};
struct int2 {
	int x, y;
};
struct alignas(16) int4 {
	int x, y, z, w;
};
struct alignas(16) uint4 {
	uint32_t x, y, z, w;
};
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline uint2 make_uint2(uint32_t x, uint32_t y) { return uint2{x, y}; }
static inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
#endif

namespace osbc {

	OSB_HDI uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
		return __float_as_uint(f);
#else
		uint32_t u;
		std::memcpy(&u, &f, 4);
		return u;
#endif
	}
	OSB_HDI float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
		return __uint_as_float(u);
#else
		float f;
		std::memcpy(&f, &u, 4);
		return f;
#endif
	}
#if defined(__CUDA_ARCH__)
	OSB_HDI float fmul(float a, float b) { return __fmul_rn(a, b); }
	OSB_HDI float fadd(float a, float b) { return __fadd_rn(a, b); }
	OSB_HDI float fsub(float a, float b) { return __fsub_rn(a, b); }
	OSB_HDI float fdiv(float a, float b) { return __fdiv_rn(a, b); }
	OSB_HDI float fsqrt(float a) { return __fsqrt_rn(a); }
	OSB_HDI int f2i_rz(float f) { return __float2int_rz(f); } // in-range inputs only
	OSB_HDI int f2i_rn(float f) { return __float2int_rn(f); }
	OSB_HDI uint32_t f2u_rz_sat(float f) { return __float2uint_rz(f); } // negative and NaN -> 0, >= 2^32 -> 0xffffffff
	OSB_HDI float i2f(int i) { return __int2float_rn(i); }
	OSB_HDI float u2flt(uint32_t i) { return __uint2float_rn(i); }
	OSB_HDI int ffs32(uint32_t x) { return __ffs((int)x) - 1; } // index of the lowest set bit, x != 0
	OSB_HDI uint32_t brev32(uint32_t x) { return __brev(x); }
	OSB_HDI int clz32(uint32_t x) { return __clz((int)x); }
	OSB_HDI uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t s) { return __byte_perm(a, b, s); }
	OSB_HDI int mulhi(int a, int b) { return __mulhi(a, b); }
	template <class T> OSB_HDI T ldg(const T *p) { return __ldg(p); }
	template <class T> OSB_HDI T ldcg(const T *p) { return __ldcg(p); } // L2 only: data another SM wrote in this kernel
	template <class T> OSB_HDI void stcs(T *p, T v) { __stcs(p, v); }
	OSB_HDI float fminf_(float a, float b) { return fminf(a, b); }
	OSB_HDI float fmaxf_(float a, float b) { return fmaxf(a, b); }
	OSB_HDI bool isnan_(float a) { return isnan(a); }
	OSB_HDI bool isinf_(float a) { return isinf(a); }
	OSB_HDI int imin(int a, int b) { return min(a, b); }
	OSB_HDI int imax(int a, int b) { return max(a, b); }
	OSB_HDI uint32_t uminu(uint32_t a, uint32_t b) { return min(a, b); }
#else
	OSB_HDI float fmul(float a, float b) { return a * b; }
	OSB_HDI float fadd(float a, float b) { return a + b; }
	OSB_HDI float fsub(float a, float b) { return a - b; }
	OSB_HDI float fdiv(float a, float b) { return a / b; }
	OSB_HDI float fsqrt(float a) { return sqrtf(a); }
	OSB_HDI int f2i_rz(float f) { return (int)f; }
	OSB_HDI int f2i_rn(float f) { return (int)lrintf(f); } // default rounding mode: nearest even
	OSB_HDI uint32_t f2u_rz_sat(float f) { return !(f > 0.f) ? 0u : (f >= 4294967296.f ? 0xffffffffu : (uint32_t)f); }
	OSB_HDI float i2f(int i) { return (float)i; }
	OSB_HDI float u2flt(uint32_t i) { return (float)i; }
	OSB_HDI int ffs32(uint32_t x) { return __builtin_ctz(x); }
	OSB_HDI uint32_t brev32(uint32_t x) {
		x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
		x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
		x = ((x >> 4) & 0x0f0f0f0fu) | ((x & 0x0f0f0f0fu) << 4);
		return __builtin_bswap32(x);
	}
	OSB_HDI int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
	OSB_HDI uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t s) {
		const uint64_t v = ((uint64_t)b << 32) | a;
		uint32_t r = 0;
		for (int i = 0; i < 4; i++) {
			const uint32_t sel = (s >> (4 * i)) & 0xf;
			uint32_t byte = (uint32_t)(v >> (8 * (sel & 7))) & 0xff;
			if (sel & 8)
				byte = (byte & 0x80) ? 0xff : 0;
			r |= byte << (8 * i);
		}
		return r;
	}
	OSB_HDI int mulhi(int a, int b) { return (int)(((int64_t)a * (int64_t)b) >> 32); }
	template <class T> OSB_HDI T ldg(const T *p) { return *p; }
	template <class T> OSB_HDI T ldcg(const T *p) { return *p; }
	template <class T> OSB_HDI void stcs(T *p, T v) { *p = v; }
	// x86 minss / maxss semantics are not needed here: callers never pass NaN (asserted by the parity runs)
	OSB_HDI float fminf_(float a, float b) { return fminf(a, b); }
	OSB_HDI float fmaxf_(float a, float b) { return fmaxf(a, b); }
	OSB_HDI bool isnan_(float a) { return std::isnan(a); }
	OSB_HDI bool isinf_(float a) { return std::isinf(a); }
	OSB_HDI int imin(int a, int b) { return a < b ? a : b; }
	OSB_HDI int imax(int a, int b) { return a > b ? a : b; }
	OSB_HDI uint32_t uminu(uint32_t a, uint32_t b) { return a < b ? a : b; }
#endif

	// truncating (C) division by 2^sh without the generic idiv
	OSB_HDI int tdiv_pow2(int x, int sh) { return (x + ((x >> 31) & ((1 << sh) - 1))) >> sh; }

} // namespace osbc
