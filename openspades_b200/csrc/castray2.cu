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

#include <algorithm>

namespace {

	// x86 cvttss2si semantics ("integer indefinite" on overflow/NaN) for (int)floorf(x), Core/Math.h:376-378
	__device__ __forceinline__ int x86_f2i(float f) {
		return (f >= -2147483648.f && f < 2147483648.f) ? __float2int_rz(f) : (int)0x80000000;
	}

	constexpr int kThreads = 128;
	constexpr int kWordsPerHit = 10;   // sizeof(osb200_rayhit) / 4
	constexpr int kCtasPerSm = 10;     // persistent grid: 48 registers x 128 threads -> 10 CTAs per SM
	constexpr int kRefillAt = 8;       // idle lanes that trigger a refill (ray set-up is the divergent part)

	struct RayState {
		float fx, fy, fz;       // distance to the next boundary per axis (fv in the reference)
		float ax, ay, az;       // |dir|
		float invX, invY, invZ; // 1 / |dir|, 0 when the component is 0
		int ix, iy, iz;         // current cell
		uint32_t sign;          // bit0..2: dir.x/y/z > 0
		uint64_t col;           // solid mask of column (ix, iy)
		int axis, steps;
		size_t ray;
	};

	__device__ __forceinline__ void write_hit(uint32_t *__restrict__ out, size_t ray, uint32_t flags, float hx, float hy,
	                                          float hz, int bx, int by, int bz, int nx, int ny, int nz) {
		uint2 *o = reinterpret_cast<uint2 *>(out + ray * kWordsPerHit); // 40-byte records are 8-byte aligned
		o[0] = make_uint2(flags, __float_as_uint(hx));
		o[1] = make_uint2(__float_as_uint(hy), __float_as_uint(hz));
		o[2] = make_uint2((uint32_t)bx, (uint32_t)by);
		o[3] = make_uint2((uint32_t)bz, (uint32_t)nx);
		o[4] = make_uint2((uint32_t)ny, (uint32_t)nz);
	}

	// IsSolidWrapped on an already fetched column (GameMap.h:91-98): z < 0 air, z >= 64 solid
	__device__ __forceinline__ bool solid_at(uint64_t col, int z) {
		if ((unsigned)z < 64u) {
			const uint32_t w = (z & 32) ? (uint32_t)(col >> 32) : (uint32_t)col;
			return (w >> (z & 31)) & 1u;
		}
		return z >= 64;
	}

	// Persistent grid (one CTA slot per resident CTA).  Every warp keeps its 32 lanes busy: a lane whose ray has
	// ended (hit, start-solid or step limit) writes its record and waits; once kRefillAt lanes wait, the warp takes
	// that many rays from the batch-wide cursor (one atomicAdd per refill) and re-arms them.  Ray lengths differ by
	// two orders of magnitude (start-solid: 0 steps, sky rays: maxSteps), so one-ray-per-thread left two thirds
	// of the lanes idle, and a static split left whole SMs idle in the tail.
	__global__ void __launch_bounds__(kThreads)
	  k_castray2(const uint64_t *__restrict__ solid, const float *__restrict__ v0s,
	             const float *__restrict__ dirs, int maxSteps, uint32_t *__restrict__ out, size_t n,
	             unsigned long long *__restrict__ cursor) {
		const int lane = threadIdx.x & 31;
		bool drained = false; // the cursor ran past n: no more refills for this warp
		RayState r;
		bool active = false;
		r.fx = r.fy = r.fz = r.ax = r.ay = r.az = r.invX = r.invY = r.invZ = 0.f;
		r.ix = r.iy = r.iz = r.axis = r.steps = 0;
		r.sign = 0;
		r.col = 0;
		r.ray = 0;

		for (;;) {
			const unsigned idle = __ballot_sync(0xffffffffu, !active);
			if (idle == 0xffffffffu && drained)
				break;
			if (!drained && (idle == 0xffffffffu || __popc(idle) >= kRefillAt)) {
				// ---- hand the idle lanes the next rays; set-up = GameMap.cpp:361-411 ----
				unsigned long long base = 0;
				if (lane == 0)
					base = atomicAdd(cursor, (unsigned long long)__popc(idle));
				base = __shfl_sync(0xffffffffu, base, 0);
				const size_t end = n;
				const size_t mine = (size_t)base + __popc(idle & ((1u << lane) - 1u));
				drained = base + __popc(idle) >= n;
				if (!active && mine < end) {
					const float v0x = v0s[mine * 3], v0y = v0s[mine * 3 + 1], v0z = v0s[mine * 3 + 2];
					float dx = dirs[mine * 3], dy = dirs[mine * 3 + 1], dz = dirs[mine * 3 + 2];
					// dir = dir.Normalize()  (Core/Math.h:369-374; GetLength = sqrtf(x*x + y*y + z*z))
					float scale = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
					if (scale != 0.f)
						scale = __fdiv_rn(1.f, scale);
					dx = __fmul_rn(dx, scale);
					dy = __fmul_rn(dy, scale);
					dz = __fmul_rn(dz, scale);
					const int ix = x86_f2i(floorf(v0x)), iy = x86_f2i(floorf(v0y)), iz = x86_f2i(floorf(v0z));
					const uint64_t col = solid[osb::tiled_index(ix & 511, iy & 511)];
					if (solid_at(col, iz)) { // startSolid (:375-384)
						write_hit(out, mine, 1u | (1u << 8), v0x, v0y, v0z, ix, iy, iz, 0, 0, 0);
					} else if (maxSteps <= 0) { // the loop never runs: hit = false, hitPos = v0, rest untouched (:503-506)
						write_hit(out, mine, 0u, v0x, v0y, v0z, 0, 0, 0, 0, 0, 0);
					} else {
						const bool px = dx > 0.f, py = dy > 0.f, pz = dz > 0.f;
						r.fx = px ? __fsub_rn((float)(ix + 1), v0x) : __fsub_rn(v0x, (float)ix);
						r.fy = py ? __fsub_rn((float)(iy + 1), v0y) : __fsub_rn(v0y, (float)iy);
						r.fz = pz ? __fsub_rn((float)(iz + 1), v0z) : __fsub_rn(v0z, (float)iz);
						r.ax = fabsf(dx);
						r.ay = fabsf(dy);
						r.az = fabsf(dz);
						r.invX = dx != 0.f ? __fdiv_rn(1.f, r.ax) : dx;
						r.invY = dy != 0.f ? __fdiv_rn(1.f, r.ay) : dy;
						r.invZ = dz != 0.f ? __fdiv_rn(1.f, r.az) : dz;
						r.sign = (px ? 1u : 0u) | (py ? 2u : 0u) | (pz ? 4u : 0u);
						r.ix = ix;
						r.iy = iy;
						r.iz = iz;
						r.col = col;
						r.steps = 0;
						r.axis = 0;
						r.ray = mine;
						active = true;
					}
				}
			}
			// ---- a few DDA steps between idle checks (:413-497) ----
#pragma unroll 1
			for (int k = 0; k < 4; k++) {
				if (active) {
					const bool hasX = r.invX != 0.f, hasY = r.invY != 0.f, hasZ = r.invZ != 0.f;
					int axis = 0;
					float t = 0.f;
					const float tx = __fmul_rn(r.fx, r.invX), ty = __fmul_rn(r.fy, r.invY), tz = __fmul_rn(r.fz, r.invZ);
					if (hasX) {
						t = tx;
						axis = 1;
					}
					if (hasY && (!axis || ty < t)) {
						t = ty;
						axis = 2;
					}
					if (hasZ && (!axis || tz < t)) {
						t = tz;
						axis = 3;
					}
					const float nfx = __fsub_rn(r.fx, __fmul_rn(r.ax, t));
					const float nfy = __fsub_rn(r.fy, __fmul_rn(r.ay, t));
					const float nfz = __fsub_rn(r.fz, __fmul_rn(r.az, t));
					r.fx = axis == 1 ? 1.f : nfx;
					r.fy = axis == 2 ? 1.f : nfy;
					r.fz = axis == 3 ? 1.f : nfz;
					r.ix += axis == 1 ? ((r.sign & 1u) ? 1 : -1) : 0;
					r.iy += axis == 2 ? ((r.sign & 2u) ? 1 : -1) : 0;
					r.iz += axis == 3 ? ((r.sign & 4u) ? 1 : -1) : 0;
					r.axis = axis;
					r.steps++;
					if (axis == 1 || axis == 2)
						r.col = solid[osb::tiled_index(r.ix & 511, r.iy & 511)];
					const bool hit = solid_at(r.col, r.iz);
					if (hit || r.steps >= maxSteps) {
						// normal = iv - next = minus the step just taken; a zero-axis step (all components 0) moves nowhere
						const int nx = axis == 1 ? ((r.sign & 1u) ? -1 : 1) : 0;
						const int ny = axis == 2 ? ((r.sign & 2u) ? -1 : 1) : 0;
						const int nz = axis == 3 ? ((r.sign & 4u) ? -1 : 1) : 0;
						float hx, hy, hz;
						if (hit) {
							hx = (r.sign & 1u) ? __fsub_rn((float)(r.ix + 1), r.fx) : __fadd_rn((float)r.ix, r.fx);
							hy = (r.sign & 2u) ? __fsub_rn((float)(r.iy + 1), r.fy) : __fadd_rn((float)r.iy, r.fy);
							hz = (r.sign & 4u) ? __fsub_rn((float)(r.iz + 1), r.fz) : __fadd_rn((float)r.iz, r.fz);
						} else { // hit = false, hitPos = v0 (:503-506); hitBlock / normal = last visited
							hx = v0s[r.ray * 3];
							hy = v0s[r.ray * 3 + 1];
							hz = v0s[r.ray * 3 + 2];
						}
						write_hit(out, r.ray, hit ? 1u : 0u, hx, hy, hz, r.ix, r.iy, r.iz, nx, ny, nz);
						active = false;
					}
				}
			}
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
	if (!c->d_rayCursor)
		OSB_CUDA(cudaMalloc(&c->d_rayCursor, sizeof(unsigned long long)));
	OSB_CUDA(cudaMemsetAsync(c->d_rayCursor, 0, sizeof(unsigned long long), c->stream));
	size_t blocks = std::min<size_t>((size_t)c->smCount * kCtasPerSm, (n + kThreads - 1) / kThreads);
	{
		osb::KLaunch kl(c, osb::K_CASTRAY2);
		k_castray2<<<(unsigned)blocks, kThreads, 0, c->stream>>>(c->d_solid, d_v0, d_dir, maxSteps,
		                                                       reinterpret_cast<uint32_t *>(d_out), n, c->d_rayCursor);
	}
	OSB_CUDA(cudaGetLastError());
	return 0;
}

// Host-buffer entry point.  The batch is cut into chunks of kPipeRays rays that flow through three stages on three
// streams -- H2D of chunk k+1, the kernel of chunk k, D2H of chunk k-1 -- over a ring of kPipeSlots device slots, so the
// call runs at the speed of the slower PCIe direction (40 B/ray out) instead of copy + kernel + copy in series.
// With pinned host buffers every copy is truly asynchronous; with pageable buffers the driver stages each copy and
// blocks the host meanwhile, so the issue order below (upload and launch chunk k+1 BEFORE downloading chunk k) is
// what keeps the GPU busy during the download.
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
	if (!c->rayPipe) {
		auto *P = new osb::RayPipe();
		c->rayPipe = P;
		OSB_CUDA(cudaStreamCreateWithFlags(&P->in, cudaStreamNonBlocking));
		OSB_CUDA(cudaStreamCreateWithFlags(&P->out, cudaStreamNonBlocking));
		for (int i = 0; i < osb::kPipeSlots; i++) {
			OSB_CUDA(cudaEventCreateWithFlags(&P->uploaded[i], cudaEventDisableTiming));
			OSB_CUDA(cudaEventCreateWithFlags(&P->computed[i], cudaEventDisableTiming));
			OSB_CUDA(cudaEventCreateWithFlags(&P->downloaded[i], cudaEventDisableTiming));
		}
	}
	osb::RayPipe *P = c->rayPipe;
	const size_t chunk = osb::kPipeRays;
	const size_t inPad = (chunk * 12 + 255) & ~(size_t)255, outPad = (chunk * sizeof(osb200_rayhit) + 255) & ~(size_t)255;
	const size_t slotBytes = 2 * inPad + outPad;
	if (c->ensure_scratch(slotBytes * osb::kPipeSlots))
		return -2;
	char *base = reinterpret_cast<char *>(c->d_scratch);
	const size_t nChunks = (n + chunk - 1) / chunk;
	// the scratch area may still be read by earlier work of the context stream
	OSB_CUDA(cudaEventRecord(P->computed[0], c->stream));
	OSB_CUDA(cudaStreamWaitEvent(P->in, P->computed[0], 0));
	auto upload = [&](size_t k) -> int {
		const int s = (int)(k % osb::kPipeSlots);
		const size_t lo = k * chunk, m = std::min(chunk, n - lo);
		char *p = base + slotBytes * s;
		if (k >= (size_t)osb::kPipeSlots) // the slot's previous results must have left the device
			OSB_CUDA(cudaStreamWaitEvent(P->in, P->downloaded[s], 0));
		OSB_CUDA(cudaMemcpyAsync(p, v0 + lo * 3, m * 12, cudaMemcpyHostToDevice, P->in));
		OSB_CUDA(cudaMemcpyAsync(p + inPad, dir + lo * 3, m * 12, cudaMemcpyHostToDevice, P->in));
		OSB_CUDA(cudaEventRecord(P->uploaded[s], P->in));
		OSB_CUDA(cudaStreamWaitEvent(c->stream, P->uploaded[s], 0));
		int rc = osb200_castray2_device(c, reinterpret_cast<float *>(p), reinterpret_cast<float *>(p + inPad), maxSteps,
		                                reinterpret_cast<osb200_rayhit *>(p + 2 * inPad), m);
		if (rc)
			return rc;
		OSB_CUDA(cudaEventRecord(P->computed[s], c->stream));
		return 0;
	};
	auto download = [&](size_t k) -> int {
		const int s = (int)(k % osb::kPipeSlots);
		const size_t lo = k * chunk, m = std::min(chunk, n - lo);
		char *p = base + slotBytes * s;
		OSB_CUDA(cudaStreamWaitEvent(P->out, P->computed[s], 0));
		OSB_CUDA(cudaMemcpyAsync(out + lo, p + 2 * inPad, m * sizeof(osb200_rayhit), cudaMemcpyDeviceToHost, P->out));
		OSB_CUDA(cudaEventRecord(P->downloaded[s], P->out));
		return 0;
	};
	if (int rc = upload(0))
		return rc;
	for (size_t k = 0; k < nChunks; k++) {
		if (k + 1 < nChunks)
			if (int rc = upload(k + 1))
				return rc;
		if (int rc = download(k))
			return rc;
	}
	OSB_CUDA(cudaStreamSynchronize(P->out));
	// later work on the context stream must not overwrite the scratch area before the last download has read it
	OSB_CUDA(cudaStreamWaitEvent(c->stream, P->downloaded[(nChunks - 1) % osb::kPipeSlots], 0));
	return 0;
}

} // extern "C"
