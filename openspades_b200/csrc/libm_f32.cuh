// Bit-exact restatements of the two libm functions the reference calls once per yaw line
// (SWMapRenderer.cpp:272 atanf in the pitch clip, :43-57 tanf in SpecialTan), so that the per-line
// set-up can run on the device and still agree with the reference running on glibc.
//
// Target: glibc 2.39 x86-64 (the image's libm; atanf/tanf there are the fdlibm-derived binary32
// routines, plain SSE scalar code without FMA).  tests/test_libm_f32.py compiles this header for
// the host and compares it against the box's libm over every float in the domains used
// (atanf: all finite floats; tanf: |x| < 1.6), bit for bit.  All arithmetic is separate
// mul/add/div in round-to-nearest; the translation unit must be built with --fmad=false.
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define OSB_HD __host__ __device__ __forceinline__
#else
#define OSB_HD static inline
#endif

namespace osbm {

	OSB_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
		return __float_as_uint(f);
#else
		uint32_t u;
		std::memcpy(&u, &f, 4);
		return u;
#endif
	}
	OSB_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
		return __uint_as_float(u);
#else
		float f;
		std::memcpy(&f, &u, 4);
		return f;
#endif
	}
	// explicit single roundings (no contraction even if a caller forgets --fmad=false)
	OSB_HD float mulf(float a, float b) {
#if defined(__CUDA_ARCH__)
		return __fmul_rn(a, b);
#else
		return a * b;
#endif
	}
	OSB_HD float addf(float a, float b) {
#if defined(__CUDA_ARCH__)
		return __fadd_rn(a, b);
#else
		return a + b;
#endif
	}
	OSB_HD float subf(float a, float b) {
#if defined(__CUDA_ARCH__)
		return __fsub_rn(a, b);
#else
		return a - b;
#endif
	}
	OSB_HD float divf(float a, float b) {
#if defined(__CUDA_ARCH__)
		return __fdiv_rn(a, b);
#else
		return a / b;
#endif
	}

	// ---- atanf: sysdeps/ieee754/flt-32/s_atanf.c ----
	OSB_HD float atanf_glibc(float x) {
		const float atanhi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f};
		const float atanlo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f};
		const float aT0 = 3.3333334327e-01f, aT1 = -2.0000000298e-01f, aT2 = 1.4285714924e-01f,
		            aT3 = -1.1111110449e-01f, aT4 = 9.0908870101e-02f, aT5 = -7.6918758452e-02f,
		            aT6 = 6.6610731184e-02f, aT7 = -5.8335702866e-02f, aT8 = 4.9768779427e-02f,
		            aT9 = -3.6531571299e-02f, aT10 = 1.6285819933e-02f;
		const int32_t hx = (int32_t)f2u(x);
		const int32_t ix = hx & 0x7fffffff;
		int id;
		if (ix >= 0x4c000000) { // |x| >= 2^25
			if (ix > 0x7f800000)
				return addf(x, x); // NaN
			if (hx > 0)
				return addf(atanhi[3], atanlo[3]);
			return subf(-atanhi[3], atanlo[3]);
		}
		if (ix < 0x3ee00000) { // |x| < 0.4375
			if (ix < 0x31000000) // |x| < 2^-29
				return x;
			id = -1;
		} else {
			x = u2f((uint32_t)ix); // fabsf
			if (ix < 0x3f980000) {     // |x| < 1.1875
				if (ix < 0x3f300000) { // 7/16 <= |x| < 11/16
					id = 0;
					x = divf(subf(addf(x, x), 1.0f), addf(2.0f, x));
				} else { // 11/16 <= |x| < 19/16
					id = 1;
					x = divf(subf(x, 1.0f), addf(x, 1.0f));
				}
			} else {
				if (ix < 0x401c0000) { // |x| < 2.4375
					id = 2;
					x = divf(subf(x, 1.5f), addf(1.0f, mulf(1.5f, x)));
				} else { // 2.4375 <= |x| < 2^25
					id = 3;
					x = divf(-1.0f, x);
				}
			}
		}
		const float z = mulf(x, x);
		const float w = mulf(z, z);
		float s1 = addf(mulf(w, aT10), aT8);
		s1 = addf(mulf(w, s1), aT6);
		s1 = addf(mulf(w, s1), aT4);
		s1 = addf(mulf(w, s1), aT2);
		s1 = addf(mulf(w, s1), aT0);
		s1 = mulf(z, s1);
		float s2 = addf(mulf(w, aT9), aT7);
		s2 = addf(mulf(w, s2), aT5);
		s2 = addf(mulf(w, s2), aT3);
		s2 = addf(mulf(w, s2), aT1);
		s2 = mulf(w, s2);
		const float xs = mulf(x, addf(s1, s2));
		if (id < 0)
			return subf(x, xs);
		const float zz = subf(atanhi[id], subf(subf(xs, atanlo[id]), x));
		return hx < 0 ? -zz : zz;
	}

	// ---- __kernel_tanf: sysdeps/ieee754/flt-32/k_tanf.c (glibc 2.39) ----
	OSB_HD float kernel_tanf(float x, float y, int iy) {
		const float pio4 = 7.8539812565e-01f, pio4lo = 3.7748947079e-08f;
		const float T0 = 3.3333334327e-01f, T1 = 1.3333334029e-01f, T2 = 5.3968254477e-02f, T3 = 2.1869488060e-02f,
		            T4 = 8.8632395491e-03f, T5 = 3.5920790397e-03f, T6 = 1.4562094584e-03f, T7 = 5.8804126456e-04f,
		            T8 = 2.4646313977e-04f, T9 = 7.8179444245e-05f, T10 = 7.1407252108e-05f,
		            T11 = -1.8558637748e-05f, T12 = 2.5907305826e-05f;
		const int32_t hx = (int32_t)f2u(x);
		const int32_t ix = hx & 0x7fffffff;
		if (ix < 0x39000000) { // |x| < 2^-13
			if ((int)x == 0) {
				if ((ix | (iy + 1)) == 0)
					return divf(1.0f, u2f((uint32_t)ix));
				else if (iy == 1)
					return x;
				else
					return divf(-1.0f, x);
			}
		}
		if (ix >= 0x3f2ca140) { // |x| >= 0.6744
			if (hx < 0) {
				x = -x;
				y = -y;
			}
			const float z0 = subf(pio4, x);
			const float w0 = subf(pio4lo, y);
			x = addf(z0, w0);
			y = 0.0f;
			if (u2f(f2u(x) & 0x7fffffffu) < 1.2207031250e-04f) // |x| < 2^-13
				return mulf((float)((1 - ((hx >> 30) & 2)) * iy), subf(1.0f, mulf((float)(2 * iy), x)));
		}
		const float z = mulf(x, x);
		float w = mulf(z, z);
		float r = addf(mulf(w, T11), T9);
		r = addf(mulf(w, r), T7);
		r = addf(mulf(w, r), T5);
		r = addf(mulf(w, r), T3);
		r = addf(mulf(w, r), T1);
		float v = addf(mulf(w, T12), T10);
		v = addf(mulf(w, v), T8);
		v = addf(mulf(w, v), T6);
		v = addf(mulf(w, v), T4);
		v = addf(mulf(w, v), T2);
		v = mulf(z, v);
		float s = mulf(z, x);
		r = addf(y, mulf(z, addf(mulf(s, addf(r, v)), y)));
		r = addf(r, mulf(T0, s));
		w = addf(x, r);
		if (ix >= 0x3f2ca140) {
			v = (float)iy;
			return mulf((float)(1 - ((hx >> 30) & 2)),
			            subf(v, mulf(2.0f, subf(x, subf(divf(mulf(w, w), addf(w, v)), r)))));
		}
		if (iy == 1)
			return w;
		// -1.0 / (x + r), accurately
		const float zt = u2f(f2u(w) & 0xfffff000u);
		v = subf(r, subf(zt, x));
		const float a = divf(-1.0f, w);
		const float t = u2f(f2u(a) & 0xfffff000u);
		s = addf(1.0f, mulf(t, zt));
		return addf(t, mulf(a, addf(s, mulf(t, v))));
	}

	// ---- tanf: sysdeps/ieee754/flt-32/s_tanf.c (glibc 2.39), valid for |x| < 120 ----
	OSB_HD float tanf_glibc(float x) {
		const int32_t hx = (int32_t)f2u(x);
		const int32_t ix = hx & 0x7fffffff;
		if (ix <= 0x3f490fda) // |x| <= pi/4
			return kernel_tanf(x, 0.0f, 1);
		if (ix > 0x7f7fffff)
			return subf(x, x); // inf / NaN -> NaN
		// reduce_fast (s_sincosf.h): n = round(x * 2/pi), y = x - n * pi/2 in binary64
		double xd = (double)x;
#if defined(__CUDA_ARCH__)
		const double r = __dmul_rn(xd, 0x1.45f306dc9c883p+23);
		const int n = ((int32_t)r + 0x800000) >> 24;
		xd = __dsub_rn(xd, __dmul_rn((double)n, 0x1.921fb54442d18p+0));
		const float y0 = __double2float_rn(xd);
		const float y1 = __double2float_rn(__dsub_rn(xd, (double)y0));
#else
		const double r = xd * 0x1.45f306dc9c883p+23;
		const int n = ((int32_t)r + 0x800000) >> 24;
		xd = xd - (double)n * 0x1.921fb54442d18p+0;
		const float y0 = (float)xd;
		const float y1 = (float)(xd - (double)y0);
#endif
		return kernel_tanf(y0, y1, 1 - ((n & 1) << 1));
	}
} // namespace osbm
