// GameMap::CastRay2 (Sources/Client/GameMap.cpp:361-507), batched: one ray per thread.
//
// Numerics: the reference is plain IEEE binary32 (sqrtf, 1/x, mul, sub, compares) compiled for
// x86-64 without FMA contraction, so this file is built with --fmad=false and uses the
// round-to-nearest intrinsics explicitly; every field of RayCastResult is reproduced bit for bit.
//
// Memory: 24 B in + 40 B out per ray (the algorithmic bytes) plus the 2 MiB
// solid bitmap, which is L2/L1-resident in 4x4-column tiles (ctx.h: tiled_index).  Result records
// are transposed through shared memory so that each warp writes its 1280 B contiguously.
#include "ctx.h"

namespace {

	// x86 cvttss2si semantics ("integer indefinite" on overflow/NaN) for (int)floorf(x), Core/Math.h:376-378
	__device__ __forceinline__ int x86_f2i(float f) {
		return (f >= -2147483648.f && f < 2147483648.f) ? __float2int_rz(f) : (int)0x80000000;
	}

	constexpr int kThreads = 256;
	constexpr int kWordsPerHit = 10; // sizeof(osb200_rayhit) / 4

	__global__ void __launch_bounds__(kThreads)
	  k_castray2(const uint64_t *__restrict__ solid, const float *__restrict__ v0s,
	             const float *__restrict__ dirs, int maxSteps, uint32_t *__restrict__ out, size_t n) {
		__shared__ uint32_t s_out[kThreads * kWordsPerHit];
		const size_t base = (size_t)blockIdx.x * kThreads;
		const size_t i = base + threadIdx.x;
		const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
		uint32_t *my = s_out + (size_t)threadIdx.x * kWordsPerHit;

		if (i < n) {
			const float v0x = v0s[i * 3], v0y = v0s[i * 3 + 1], v0z = v0s[i * 3 + 2];
			float dx = dirs[i * 3], dy = dirs[i * 3 + 1], dz = dirs[i * 3 + 2];

			// dir = dir.Normalize()  (Core/Math.h:369-374; GetLength = sqrtf(x*x + y*y + z*z))
			float scale = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
			if (scale != 0.f)
				scale = __fdiv_rn(1.f, scale);
			dx = __fmul_rn(dx, scale);
			dy = __fmul_rn(dy, scale);
			dz = __fmul_rn(dz, scale);

			int ix = x86_f2i(floorf(v0x)), iy = x86_f2i(floorf(v0y)), iz = x86_f2i(floorf(v0z));

			uint32_t flags = 0; // hit | startSolid << 8
			float hx = v0x, hy = v0y, hz = v0z;
			int bx = 0, by = 0, bz = 0, nx = 0, ny = 0, nz = 0;

			// IsSolidWrapped (GameMap.h:91-98): z < 0 air, z >= 64 solid, x/y wrap & 511
			uint64_t col = solid[osb::tiled_index(ix & 511, iy & 511)];
			bool startSolid = iz >= 64 || (iz >= 0 && ((col >> iz) & 1ull));
			if (startSolid) {
				flags = 1u | (1u << 8);
				bx = ix;
				by = iy;
				bz = iz;
			} else {
				const bool px = dx > 0.f, py = dy > 0.f, pz = dz > 0.f;
				float fx = px ? __fsub_rn((float)(ix + 1), v0x) : __fsub_rn(v0x, (float)ix);
				float fy = py ? __fsub_rn((float)(iy + 1), v0y) : __fsub_rn(v0y, (float)iy);
				float fz = pz ? __fsub_rn((float)(iz + 1), v0z) : __fsub_rn(v0z, (float)iz);
				const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
				const float invX = dx != 0.f ? __fdiv_rn(1.f, ax) : dx;
				const float invY = dy != 0.f ? __fdiv_rn(1.f, ay) : dy;
				const float invZ = dz != 0.f ? __fdiv_rn(1.f, az) : dz;
				const bool hasX = invX != 0.f, hasY = invY != 0.f, hasZ = invZ != 0.f;
				const int sx = px ? 1 : -1, sy = py ? 1 : -1, sz = pz ? 1 : -1;

				for (int s = 0; s < maxSteps; s++) {
					int axis = 0;
					float t = 0.f;
					if (hasX) {
						t = __fmul_rn(fx, invX);
						axis = 1;
					}
					if (hasY) {
						float ty = __fmul_rn(fy, invY);
						if (!axis || ty < t) {
							t = ty;
							axis = 2;
						}
					}
					if (hasZ) {
						float tz = __fmul_rn(fz, invZ);
						if (!axis || tz < t) {
							t = tz;
							axis = 3;
						}
					}
					fx = axis == 1 ? 1.f : __fsub_rn(fx, __fmul_rn(ax, t));
					fy = axis == 2 ? 1.f : __fsub_rn(fy, __fmul_rn(ay, t));
					fz = axis == 3 ? 1.f : __fsub_rn(fz, __fmul_rn(az, t));

					int qx = ix + (axis == 1 ? sx : 0);
					int qy = iy + (axis == 2 ? sy : 0);
					int qz = iz + (axis == 3 ? sz : 0);
					bx = qx;
					by = qy;
					bz = qz;
					nx = ix - qx;
					ny = iy - qy;
					nz = iz - qz;
					if (axis == 1 || axis == 2)
						col = solid[osb::tiled_index(qx & 511, qy & 511)];
					bool isSolid = qz >= 64 || (qz >= 0 && ((col >> qz) & 1ull));
					if (isSolid) {
						hx = px ? __fsub_rn((float)(qx + 1), fx) : __fadd_rn((float)qx, fx);
						hy = py ? __fsub_rn((float)(qy + 1), fy) : __fadd_rn((float)qy, fy);
						hz = pz ? __fsub_rn((float)(qz + 1), fz) : __fadd_rn((float)qz, fz);
						flags = 1u;
						break;
					}
					ix = qx;
					iy = qy;
					iz = qz;
				}
			}
			my[0] = flags;
			my[1] = __float_as_uint(hx);
			my[2] = __float_as_uint(hy);
			my[3] = __float_as_uint(hz);
			my[4] = (uint32_t)bx;
			my[5] = (uint32_t)by;
			my[6] = (uint32_t)bz;
			my[7] = (uint32_t)nx;
			my[8] = (uint32_t)ny;
			my[9] = (uint32_t)nz;
		}
		__syncwarp();
		// coalesced write-out of this warp's 32 records (320 words)
		const size_t wbase = base + (size_t)warp * 32;
		if (wbase < n) {
			size_t cnt = n - wbase;
			if (cnt > 32)
				cnt = 32;
			const uint32_t *src = s_out + (size_t)warp * 32 * kWordsPerHit;
			uint32_t *dst = out + wbase * kWordsPerHit;
			const int words = (int)cnt * kWordsPerHit;
			for (int w = lane; w < words; w += 32)
				dst[w] = src[w];
		}
	}
} // namespace

extern "C" {

int osb200_castray2_device(osb200_ctx *c, const float *d_v0, const float *d_dir, int maxSteps,
                           osb200_rayhit *d_out, size_t n) {
	if (!c || !c->mapLoaded) {
		osb::set_error("osb200_castray2: no context or no map uploaded");
		return -1;
	}
	if (n == 0)
		return 0;
	if (!d_v0 || !d_dir || !d_out) {
		osb::set_error("osb200_castray2: null buffer");
		return -1;
	}
	static_assert(sizeof(osb200_rayhit) == 40, "RayCastResult layout");
	OSB_CUDA(cudaSetDevice(c->device));
	size_t blocks = (n + kThreads - 1) / kThreads;
	{
		osb::KLaunch kl(c, osb::K_CASTRAY2);
		k_castray2<<<(unsigned)blocks, kThreads, 0, c->stream>>>(c->d_solid, d_v0, d_dir, maxSteps,
		                                                       reinterpret_cast<uint32_t *>(d_out), n);
	}
	OSB_CUDA(cudaGetLastError());
	return 0;
}

int osb200_castray2(osb200_ctx *c, const float *v0, const float *dir, int maxSteps, osb200_rayhit *out,
                    size_t n) {
	if (!c || !c->mapLoaded) {
		osb::set_error("osb200_castray2: no context or no map uploaded");
		return -1;
	}
	if (n == 0)
		return 0;
	if (!v0 || !dir || !out) {
		osb::set_error("osb200_castray2: null buffer");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	const size_t inB = n * 12, outB = n * sizeof(osb200_rayhit);
	const size_t inPad = (inB + 255) & ~(size_t)255;
	if (c->ensure_scratch(2 * inPad + outB))
		return -2;
	char *p = reinterpret_cast<char *>(c->d_scratch);
	float *dv = reinterpret_cast<float *>(p), *dd = reinterpret_cast<float *>(p + inPad);
	auto *dout = reinterpret_cast<osb200_rayhit *>(p + 2 * inPad);
	OSB_CUDA(cudaMemcpyAsync(dv, v0, inB, cudaMemcpyHostToDevice, c->stream));
	OSB_CUDA(cudaMemcpyAsync(dd, dir, inB, cudaMemcpyHostToDevice, c->stream));
	int rc = osb200_castray2_device(c, dv, dd, maxSteps, dout, n);
	if (rc)
		return rc;
	OSB_CUDA(cudaMemcpyAsync(out, dout, outB, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

} // extern "C"
