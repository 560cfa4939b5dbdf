// Device-side equivalents of the reference's numeric helpers in the three variants it has
// (Sources/Draw/SWFeatureLevel.h:89-121):
//   IEEE : fastRcp = 1/x, fastRSqrt = 1/sqrtf(x), fastDiv = a/b            (build without SSE)
//   SSE* : RCPSS / RSQRTSS approximations, reproduced from tables calibrated on the host CPU
//          (sse_calibrate.cpp documents and verifies the table model).
#pragma once
#include <cstdint>

#include "hd.cuh"

namespace osbn {

	struct SseTables {
		const uint32_t *rcp;   // 1 << rcpBits entries
		const uint32_t *rsqrt; // 2 << rsqrtBits entries
		int rcpBits, rsqrtBits;
	};

	// x86 cvttss2si: out-of-range and NaN give the "integer indefinite" value
	OSB_HDI int x86_f2i(float f) {
		return (f >= -2147483648.f && f < 2147483648.f) ? osbc::f2i_rz(f) : (int)0x80000000;
	}
	// x86 cvtps2dq under the default MXCSR (round to nearest even)
	OSB_HDI int x86_f2i_rn(float f) {
		return (f >= -2147483648.f && f < 2147483648.f) ? osbc::f2i_rn(f) : (int)0x80000000;
	}

	OSB_HDI float rcpss(float x, const SseTables &t) {
		const uint32_t u = osbc::f2u(x);
		const uint32_t sign = u & 0x80000000u, e = (u >> 23) & 0xff, m = u & 0x7fffff;
		if (e - 1u < 252u) {
			// normal input, normal result (table entries lie in [0.5, 1]: biased exponent 126 or 127, so the
			// result exponent 127 - e + {126,127} stays in 1..253): add the exponent difference in place
			const uint32_t r = osbc::ldg(t.rcp + (m >> (23 - t.rcpBits)));
			return osbc::u2f(sign | (r + ((127u - e) << 23)));
		}
		if (e == 0)
			return osbc::u2f(sign | 0x7f800000u); // zero / denormal -> +-inf
		if (e == 255)
			return osbc::u2f(m ? (u | 0x00400000u) : sign);
		const uint32_t r = osbc::ldg(t.rcp + (m >> (23 - t.rcpBits)));
		const int re = (int)((r >> 23) & 0xff) - ((int)e - 127);
		if (re <= 0)
			return osbc::u2f(sign); // underflow flushes to zero
		return osbc::u2f(sign | ((uint32_t)re << 23) | (r & 0x7fffff));
	}

	OSB_HDI float rsqrtss(float x, const SseTables &t) {
		const uint32_t u = osbc::f2u(x);
		const uint32_t sign = u & 0x80000000u, e = (u >> 23) & 0xff, m = u & 0x7fffff;
		if (e == 255 && m)
			return osbc::u2f(u | 0x00400000u);
		if (e == 0)
			return osbc::u2f(sign | 0x7f800000u);
		if (sign)
			return osbc::u2f(0xffc00000u);
		if (e == 255)
			return 0.f;
		const int E = (int)e - 127, p = E & 1;
		const uint32_t r = osbc::ldg(t.rsqrt + (((uint32_t)p << t.rsqrtBits) | (m >> (23 - t.rsqrtBits))));
		const int re = (int)((r >> 23) & 0xff) - ((E - p) >> 1);
		return osbc::u2f(((uint32_t)re << 23) | (r & 0x7fffff));
	}

	// rcpss for callers that only use the value when x is a normal number of magnitude >= 2^-126 (ToSpecialTan uses
	// it for |x| > 1, the DDA for distances >= 1e-20): normal input, normal result, no branch.  Huge / inf / NaN
	// inputs still go through the general routine; for zero and denormals the value is unspecified.
	OSB_HDI float rcpss_normal(float x, const SseTables &t) {
		const uint32_t u = osbc::f2u(x);
		const uint32_t e = (u >> 23) & 0xff;
		const uint32_t r = osbc::ldg(t.rcp + ((u & 0x7fffff) >> (23 - t.rcpBits)));
		if (e >= 253u)
			return rcpss(x, t);
		return osbc::u2f((u & 0x80000000u) | (r + ((127u - e) << 23)));
	}

	template <bool kSse> struct Num {
		SseTables t;
		// reciprocal whose value is only used when |b| > 1 (or b a DDA distance >= 1e-20): skips nothing in IEEE
		// mode, takes the branch-free table path in SSE mode
		OSB_HDM float rcpBig(float b) const {
			if (kSse)
				return rcpss_normal(b, t);
			return osbc::fdiv(1.f, b);
		}
		// Hot-loop reciprocals of the line walker: the argument is a normal number with biased exponent in 1..252
		// (a DDA distance in [1e-20, 256), or |invDist * (z - cz)| in (1, 64e20]), so the table path needs neither
		// the special-value nor the range test: result bits = table + (127 - e) << 23, sign copied.
		OSB_HDM float rcpHot(float b) const {
			if (kSse) {
				const uint32_t u = osbc::f2u(b);
				const uint32_t r = osbc::ldg(t.rcp + ((u & 0x7fffffu) >> (23 - t.rcpBits)));
				return osbc::u2f((r + 0x3f800000u - (u & 0x7f800000u)) | (u & 0x80000000u));
			}
			return osbc::fdiv(1.f, b);
		}
		// ToSpecialTan for |v| > 1 (SWMapRenderer.cpp:59-66): -2 - rcp(v) for v < -1, 2 - rcp(v) for v > 1.  Both are
		// sign(v) * (2 - rcp(|v|)) -- the reciprocal (table or division) and the subtraction are odd functions under
		// round-to-nearest -- so the sign is dealt with once, by copying it onto the positive result.
		OSB_HDM float specialTanTail(float v) const {
			const uint32_t u = osbc::f2u(v);
			float ra;
			if (kSse) {
				const uint32_t r = t.rcp[(u & 0x7fffffu) >> (23 - t.rcpBits)]; // (plain load: k_walk_lines keeps the table in shared memory)
				ra = osbc::u2f(r + 0x3f800000u - (u & 0x7f800000u));
			} else {
				ra = osbc::fdiv(1.f, osbc::u2f(u & 0x7fffffffu));
			}
			return osbc::u2f(osbc::f2u(osbc::fsub(2.f, ra)) | (u & 0x80000000u));
		}
		OSB_HDM float rcpHotPos(float b) const { // b > 0
			if (kSse) {
				const uint32_t u = osbc::f2u(b);
				const uint32_t r = t.rcp[(u & 0x7fffffu) >> (23 - t.rcpBits)];
				return osbc::u2f(r + 0x3f800000u - (u & 0x7f800000u));
			}
			return osbc::fdiv(1.f, b);
		}
		OSB_HDM float rcp(float b) const {
			if (kSse)
				return rcpss(b, t);
			return osbc::fdiv(1.f, b);
		}
		OSB_HDM float div(float a, float b) const {
			if (kSse)
				return osbc::fmul(a, rcpss(b, t));
			return osbc::fdiv(a, b);
		}
		OSB_HDM float rsqrt(float b) const {
			if (kSse)
				return rsqrtss(b, t);
			return osbc::fdiv(1.f, osbc::fsqrt(b));
		}
		OSB_HDM float sqrt(float s) const { // fastSqrt, SWFeatureLevel.h:117-121
			if (s == 0.f)
				return 0.f;
			return osbc::fmul(s, rsqrt(s));
		}
		// ToSpecialTan, SWMapRenderer.cpp:59-66
		OSB_HDM float toSpecialTan(float v) const {
			if (v < -1.f)
				return osbc::fsub(-2.f, rcp(v));
			else if (v > 1.f)
				return osbc::fsub(2.f, rcp(v));
			return v;
		}
	};
} // namespace osbn
