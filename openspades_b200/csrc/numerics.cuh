// Device-side equivalents of the reference's numeric helpers in the three variants it has
// (Sources/Draw/SWFeatureLevel.h:89-121):
//   IEEE : fastRcp = 1/x, fastRSqrt = 1/sqrtf(x), fastDiv = a/b            (build without SSE)
//   SSE* : RCPSS / RSQRTSS approximations, reproduced from tables calibrated on the host CPU
//          (sse_calibrate.cpp documents and verifies the table model).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace osbn {

	struct SseTables {
		const uint32_t *rcp;   // 1 << rcpBits entries
		const uint32_t *rsqrt; // 2 << rsqrtBits entries
		int rcpBits, rsqrtBits;
	};

	// x86 cvttss2si: out-of-range and NaN give the "integer indefinite" value
	__device__ __forceinline__ int x86_f2i(float f) {
		return (f >= -2147483648.f && f < 2147483648.f) ? __float2int_rz(f) : (int)0x80000000;
	}
	// x86 cvtps2dq under the default MXCSR (round to nearest even)
	__device__ __forceinline__ int x86_f2i_rn(float f) {
		return (f >= -2147483648.f && f < 2147483648.f) ? __float2int_rn(f) : (int)0x80000000;
	}

	__device__ __forceinline__ float rcpss(float x, const SseTables &t) {
		const uint32_t u = __float_as_uint(x);
		const uint32_t sign = u & 0x80000000u, e = (u >> 23) & 0xff, m = u & 0x7fffff;
		if (e - 1u < 252u) {
			// normal input, normal result (table entries lie in [0.5, 1]: biased exponent 126 or 127, so the
			// result exponent 127 - e + {126,127} stays in 1..253): add the exponent difference in place
			const uint32_t r = __ldg(t.rcp + (m >> (23 - t.rcpBits)));
			return __uint_as_float(sign | (r + ((127u - e) << 23)));
		}
		if (e == 0)
			return __uint_as_float(sign | 0x7f800000u); // zero / denormal -> +-inf
		if (e == 255)
			return __uint_as_float(m ? (u | 0x00400000u) : sign);
		const uint32_t r = __ldg(t.rcp + (m >> (23 - t.rcpBits)));
		const int re = (int)((r >> 23) & 0xff) - ((int)e - 127);
		if (re <= 0)
			return __uint_as_float(sign); // underflow flushes to zero
		return __uint_as_float(sign | ((uint32_t)re << 23) | (r & 0x7fffff));
	}

	__device__ __forceinline__ float rsqrtss(float x, const SseTables &t) {
		const uint32_t u = __float_as_uint(x);
		const uint32_t sign = u & 0x80000000u, e = (u >> 23) & 0xff, m = u & 0x7fffff;
		if (e == 255 && m)
			return __uint_as_float(u | 0x00400000u);
		if (e == 0)
			return __uint_as_float(sign | 0x7f800000u);
		if (sign)
			return __uint_as_float(0xffc00000u);
		if (e == 255)
			return 0.f;
		const int E = (int)e - 127, p = E & 1;
		const uint32_t r = __ldg(t.rsqrt + (((uint32_t)p << t.rsqrtBits) | (m >> (23 - t.rsqrtBits))));
		const int re = (int)((r >> 23) & 0xff) - ((E - p) >> 1);
		return __uint_as_float(((uint32_t)re << 23) | (r & 0x7fffff));
	}

	// rcpss for callers that only use the value when x is a normal number of magnitude >= 2^-126 (ToSpecialTan uses
	// it for |x| > 1, the DDA for distances >= 1e-20): normal input, normal result, no branch.  Huge / inf / NaN
	// inputs still go through the general routine; for zero and denormals the value is unspecified.
	__device__ __forceinline__ float rcpss_normal(float x, const SseTables &t) {
		const uint32_t u = __float_as_uint(x);
		const uint32_t e = (u >> 23) & 0xff;
		const uint32_t r = __ldg(t.rcp + ((u & 0x7fffff) >> (23 - t.rcpBits)));
		if (e >= 253u)
			return rcpss(x, t);
		return __uint_as_float((u & 0x80000000u) | (r + ((127u - e) << 23)));
	}

	template <bool kSse> struct Num {
		SseTables t;
		// reciprocal whose value is only used when |b| > 1 (or b a DDA distance >= 1e-20): skips nothing in IEEE
		// mode, takes the branch-free table path in SSE mode
		__device__ __forceinline__ float rcpBig(float b) const {
			if (kSse)
				return rcpss_normal(b, t);
			return __fdiv_rn(1.f, b);
		}
		__device__ __forceinline__ float rcp(float b) const {
			if (kSse)
				return rcpss(b, t);
			return __fdiv_rn(1.f, b);
		}
		__device__ __forceinline__ float div(float a, float b) const {
			if (kSse)
				return __fmul_rn(a, rcpss(b, t));
			return __fdiv_rn(a, b);
		}
		__device__ __forceinline__ float rsqrt(float b) const {
			if (kSse)
				return rsqrtss(b, t);
			return __fdiv_rn(1.f, __fsqrt_rn(b));
		}
		__device__ __forceinline__ float sqrt(float s) const { // fastSqrt, SWFeatureLevel.h:117-121
			if (s == 0.f)
				return 0.f;
			return __fmul_rn(s, rsqrt(s));
		}
		// ToSpecialTan, SWMapRenderer.cpp:59-66
		__device__ __forceinline__ float toSpecialTan(float v) const {
			if (v < -1.f)
				return __fsub_rn(-2.f, rcp(v));
			else if (v > 1.f)
				return __fsub_rn(2.f, rcp(v));
			return v;
		}
	};
} // namespace osbn
