// Host-side calibration of the x86 RCPSS / RSQRTSS approximation instructions.
//
// The reference's default build evaluates fastRcp / fastRSqrt / fastDiv with these instructions
// (Sources/Draw/SWFeatureLevel.h:89-116).  Their results are implementation-defined (Intel and
// AMD differ), so to reproduce "the reference SWRenderer on this host" bit for bit the device
// kernels look the mantissa up in tables measured here, on the CPU the reference would run on.
// Both instructions are (on every implementation we have seen) a function of the top k mantissa
// bits (and exponent parity for rsqrt) with the exponent handled exactly; calibration finds k,
// verifies that model exhaustively over all 2^23 mantissas, and fails if it does not hold.
//
// This is not a CPU fallback of any compute path: nothing here renders or casts rays.
#include <cstdint>
#include <cstring>

#include "../../include/osb200.h"

#if defined(__x86_64__) || defined(__i386__)
#include <xmmintrin.h>
#define OSB_HAVE_SSE 1
#else
#define OSB_HAVE_SSE 0
#endif

namespace osb {
	void set_error(const char *fmt, ...);
}

#if OSB_HAVE_SSE
static inline uint32_t f2u(float f) {
	uint32_t u;
	std::memcpy(&u, &f, 4);
	return u;
}
static inline float u2f(uint32_t u) {
	float f;
	std::memcpy(&f, &u, 4);
	return f;
}
static inline uint32_t hw_rcp(uint32_t u) { return f2u(_mm_cvtss_f32(_mm_rcp_ss(_mm_set_ss(u2f(u))))); }
static inline uint32_t hw_rsqrt(uint32_t u) { return f2u(_mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(u2f(u))))); }

// smallest k such that f is constant over every block of 2^(23-k) consecutive mantissas
template <class F> static int find_bits(F f, uint32_t expBits, int maxBits) {
	for (int k = 1; k <= maxBits; k++) {
		const uint32_t blk = 1u << (23 - k);
		bool ok = true;
		for (uint32_t m = 0; m < (1u << 23) && ok; m += blk) {
			const uint32_t r0 = f(expBits | m);
			for (uint32_t j = 1; j < blk; j++)
				if (f(expBits | (m + j)) != r0) {
					ok = false;
					break;
				}
		}
		if (ok)
			return k;
	}
	return -1;
}
#endif

// Software model shared with the device code (numerics.cuh implements the same arithmetic).
static uint32_t model_rcp(uint32_t u, const uint32_t *tab, int bits) {
	const uint32_t sign = u & 0x80000000u, e = (u >> 23) & 0xff, m = u & 0x7fffff;
	if (e == 0)
		return sign | 0x7f800000u;
	if (e == 255)
		return m ? (u | 0x00400000u) : sign;
	const uint32_t t = tab[m >> (23 - bits)];
	const int re = (int)((t >> 23) & 0xff) - ((int)e - 127);
	if (re <= 0)
		return sign;
	return sign | ((uint32_t)re << 23) | (t & 0x7fffff);
}
static uint32_t model_rsqrt(uint32_t u, const uint32_t *tab, int bits) {
	const uint32_t sign = u & 0x80000000u, e = (u >> 23) & 0xff, m = u & 0x7fffff;
	if (e == 255 && m)
		return u | 0x00400000u;
	if (e == 0)
		return sign | 0x7f800000u;
	if (sign)
		return 0xffc00000u;
	if (e == 255)
		return 0;
	const int E = (int)e - 127, p = E & 1;
	const uint32_t t = tab[((uint32_t)p << bits) | (m >> (23 - bits))];
	const int re = (int)((t >> 23) & 0xff) - ((E - p) >> 1);
	return ((uint32_t)re << 23) | (t & 0x7fffff);
}

extern "C" OSB200_API int osb200_calibrate_sse(uint32_t *rcp_table, int *rcp_bits, uint32_t *rsqrt_table,
                                               int *rsqrt_bits, int max_bits) {
#if !OSB_HAVE_SSE
	osb::set_error("osb200_calibrate_sse: host CPU has no RCPSS/RSQRTSS (not x86)");
	return -4;
#else
	if (!rcp_table || !rcp_bits || !rsqrt_table || !rsqrt_bits || max_bits < 1 || max_bits > 16) {
		osb::set_error("osb200_calibrate_sse: bad argument");
		return -1;
	}
	const int kr = find_bits(hw_rcp, 0x3f800000u, max_bits);
	int ks0 = find_bits(hw_rsqrt, 0x3f800000u, max_bits);
	int ks1 = find_bits(hw_rsqrt, 0x40000000u, max_bits);
	const int ks = ks0 > ks1 ? ks0 : ks1;
	if (kr < 0 || ks0 < 0 || ks1 < 0) {
		osb::set_error("osb200_calibrate_sse: RCPSS/RSQRTSS on this CPU are not piecewise constant within "
		               "%d mantissa bits; SSE numerics cannot be emulated (use OSB200_NUM_IEEE)", max_bits);
		return -4;
	}
	for (uint32_t i = 0; i < (1u << kr); i++)
		rcp_table[i] = hw_rcp(0x3f800000u | (i << (23 - kr)));
	for (uint32_t i = 0; i < (1u << ks); i++) {
		rsqrt_table[i] = hw_rsqrt(0x3f800000u | (i << (23 - ks)));
		rsqrt_table[(1u << ks) | i] = hw_rsqrt(0x40000000u | (i << (23 - ks)));
	}
	// verify the exponent model on a spread of exponents and the special values
	uint32_t rng = 12345;
	for (int it = 0; it < 2000000; it++) {
		rng = rng * 1664525u + 1013904223u;
		uint32_t u = rng ^ (rng >> 7) ^ ((uint32_t)it << 20);
		if (model_rcp(u, rcp_table, kr) != hw_rcp(u) && ((u >> 23) & 0xff) != 255) {
			osb::set_error("osb200_calibrate_sse: RCPSS model mismatch at %08x", u);
			return -4;
		}
		uint32_t a = hw_rsqrt(u), b = model_rsqrt(u, rsqrt_table, ks);
		bool nanBoth = ((a & 0x7f800000u) == 0x7f800000u && (a & 0x7fffff)) &&
		               ((b & 0x7f800000u) == 0x7f800000u && (b & 0x7fffff));
		if (a != b && !nanBoth) {
			osb::set_error("osb200_calibrate_sse: RSQRTSS model mismatch at %08x", u);
			return -4;
		}
	}
	*rcp_bits = kr;
	*rsqrt_bits = ks;
	return 0;
#endif
}
