// GameMap::CastRay2 (Sources/Client/GameMap.cpp:361-507), batched: one ray per thread.
//
// Numerics: the reference is plain IEEE binary32 (sqrtf, 1/x, mul, sub, compares) compiled for
// x86-64 without FMA contraction, so this file is built with --fmad=false and uses the
// round-to-nearest intrinsics explicitly; every field of RayCastResult is reproduced bit for bit.
//
// Memory: 24 B in + 40 B out per ray (the algorithmic bytes) plus the 2 MiB
// solid bitmap, which is L2/L1-resident in 4x4-column tiles (ctx.h: tiled_index).  Each lane writes its own
// 40-byte record as five 8-byte stores when its ray ends (rays of a warp end at different steps, so there is no
// common point at which a warp-wide transposed write could happen; L2 merges the sectors before they reach HBM).
#include "ctx.h"

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

	// x86 cvttss2si semantics ("integer indefinite" on overflow/NaN) for (int)floorf(x), Core/Math.h:376-378
	__device__ __forceinline__ int x86_f2i(float f) {
		return (f >= -2147483648.f && f < 2147483648.f) ? __float2int_rz(f) : (int)0x80000000;
	}

	// column of the map copy the ray kernel reads: GameMap's own [x][y] order, wrapped (measured against the renderer's
	// 4x4 tiles: 5.62 against 5.99 ms per 16 Mi rays -- three address instructions instead of eight in an ALU-bound step)
	__device__ __forceinline__ uint32_t ray_index(int x, int y) { return ((uint32_t)(x & 511) << 9) | (uint32_t)(y & 511); }
	__global__ void k_ray_map_xy(const uint64_t *__restrict__ solidT, uint64_t *__restrict__ solidXY) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x; // [x][y]
		solidXY[i] = solidT[osb::tiled_index(i >> 9, i & 511)];
	}

	constexpr int kThreads = 128;
	constexpr int kWordsPerHit = 10;   // sizeof(osb200_rayhit) / 4
	constexpr int kCtasPerSm = 10;     // persistent grid: 48 registers x 128 threads -> 10 CTAs per SM
	constexpr int kRefillAt = 8;       // idle lanes that trigger a refill (ray set-up is the divergent part)
#ifndef OSB_RAY_STEPS_PER_CHECK
#define OSB_RAY_STEPS_PER_CHECK 8
#endif
	constexpr int kStepsPerCheck = OSB_RAY_STEPS_PER_CHECK; // DDA steps between two looks at the warp's idle lanes

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

	// The reference's loop as written (GameMap.cpp:413-497), one ray from start to end by one lane.  Only rays whose
	// normalised direction holds a NaN come here (an infinite or NaN input component): the fast step below folds the
	// "axis present" tests into the comparisons, which orders NaNs differently.  Garbage in, the reference's garbage out.
	__device__ __noinline__ void cast_ray_generic(const uint64_t *__restrict__ solid, uint32_t *__restrict__ out, size_t ray,
	                                              float v0x, float v0y, float v0z, float dx, float dy, float dz, int ix, int iy,
	                                              int iz, uint64_t col, int maxSteps) {
		const bool px = dx > 0.f, py = dy > 0.f, pz = dz > 0.f;
		float fx = px ? __fsub_rn((float)(ix + 1), v0x) : __fsub_rn(v0x, (float)ix);
		float fy = py ? __fsub_rn((float)(iy + 1), v0y) : __fsub_rn(v0y, (float)iy);
		float fz = pz ? __fsub_rn((float)(iz + 1), v0z) : __fsub_rn(v0z, (float)iz);
		const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
		const float invX = dx != 0.f ? __fdiv_rn(1.f, ax) : dx;
		const float invY = dy != 0.f ? __fdiv_rn(1.f, ay) : dy;
		const float invZ = dz != 0.f ? __fdiv_rn(1.f, az) : dz;
		const bool hasX = invX != 0.f, hasY = invY != 0.f, hasZ = invZ != 0.f;
		int axis = 0;
		bool hit = false;
		for (int steps = 0; steps < maxSteps && !hit; steps++) {
			axis = 0;
			float t = 0.f;
			const float tx = __fmul_rn(fx, invX), ty = __fmul_rn(fy, invY), tz = __fmul_rn(fz, invZ);
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
			const float nfx = __fsub_rn(fx, __fmul_rn(ax, t)), nfy = __fsub_rn(fy, __fmul_rn(ay, t)),
			            nfz = __fsub_rn(fz, __fmul_rn(az, t));
			fx = axis == 1 ? 1.f : nfx;
			fy = axis == 2 ? 1.f : nfy;
			fz = axis == 3 ? 1.f : nfz;
			ix += axis == 1 ? (px ? 1 : -1) : 0;
			iy += axis == 2 ? (py ? 1 : -1) : 0;
			iz += axis == 3 ? (pz ? 1 : -1) : 0;
			if (axis == 1 || axis == 2)
				col = solid[ray_index(ix, iy)];
			hit = solid_at(col, iz);
		}
		const int nx = axis == 1 ? (px ? -1 : 1) : 0, ny = axis == 2 ? (py ? -1 : 1) : 0, nz = axis == 3 ? (pz ? -1 : 1) : 0;
		float hx = v0x, hy = v0y, hz = v0z;
		if (hit) {
			hx = px ? __fsub_rn((float)(ix + 1), fx) : __fadd_rn((float)ix, fx);
			hy = py ? __fsub_rn((float)(iy + 1), fy) : __fadd_rn((float)iy, fy);
			hz = pz ? __fsub_rn((float)(iz + 1), fz) : __fadd_rn((float)iz, fz);
		}
		write_hit(out, ray, hit ? 1u : 0u, hx, hy, hz, ix, iy, iz, nx, ny, nz);
	}

	struct RayState {
		float fx, fy, fz;       // distance to the next boundary per axis (fv in the reference)
		float ax, ay, az;       // |dir|
		float invX, invY, invZ; // 1 / |dir|, 0 when the component is 0
		float bX, bY, bZ;       // 0, or +inf for an axis the ray never crosses (its t must lose every comparison)
		int ix, iy, iz;         // current cell
		uint32_t sign;          // bit0..2: dir.x/y/z > 0
		uint64_t col;           // solid mask of column (ix, iy)
		int left;               // steps the ray may still take; 0 = this lane has no ray
		size_t ray;
	};

	// Persistent grid (one CTA slot per resident CTA).  Every warp keeps its 32 lanes busy: a lane whose ray has
	// ended (hit, start-solid or step limit) writes its record and waits; once kRefillAt lanes wait, the warp takes
	// that many rays from the batch-wide cursor (one atomicAdd per refill) and re-arms them.  Ray lengths differ by
	// two orders of magnitude (start-solid: 0 steps, sky rays: maxSteps), so one-ray-per-thread left two thirds
	// of the lanes idle, and a static split left whole SMs idle in the tail.
	//
	// The step is bound by the ALU pipe (compares, selects, logic: ncu 84 % of its peak in round 1), so it is written
	// to need few of those: the reference's "is this axis crossed at all" tests (hasX/Y/Z, :417-419,432-446) are folded
	// into the t of the axis -- t = fma(f, inv, bias) with bias = +inf for an axis whose direction component is 0
	// (the product is exact-rounded once and adding 0 does not change it), which then loses both "<" comparisons
	// exactly where the reference skips the axis; a ray with no axis at all never moves and is answered at set-up.
	__global__ void __launch_bounds__(kThreads, kCtasPerSm)
	  k_castray2(const uint64_t *__restrict__ solid, const float *__restrict__ v0s,
	             const float *__restrict__ dirs, int maxSteps, uint32_t *__restrict__ out, size_t n,
	             unsigned long long *__restrict__ cursor) {
		const int lane = threadIdx.x & 31;
		bool drained = false; // the cursor ran past n: no more refills for this warp
		RayState r;
		r.fx = r.fy = r.fz = r.ax = r.ay = r.az = r.invX = r.invY = r.invZ = r.bX = r.bY = r.bZ = 0.f;
		r.ix = r.iy = r.iz = 0;
		r.left = 0;
		r.sign = 0;
		r.col = 0;
		r.ray = 0;
		const float kInf = __int_as_float(0x7f800000);

		for (;;) {
			const unsigned idle = __ballot_sync(0xffffffffu, r.left == 0);
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
				if (r.left == 0 && mine < end) {
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
					const uint64_t col = solid[ray_index(ix, iy)];
					if (solid_at(col, iz)) { // startSolid (:375-384)
						write_hit(out, mine, 1u | (1u << 8), v0x, v0y, v0z, ix, iy, iz, 0, 0, 0);
					} else if (maxSteps <= 0) { // the loop never runs: hit = false, hitPos = v0, rest untouched (:503-506)
						write_hit(out, mine, 0u, v0x, v0y, v0z, 0, 0, 0, 0, 0, 0);
					} else if (dx != dx || dy != dy || dz != dz) {
						cast_ray_generic(solid, out, mine, v0x, v0y, v0z, dx, dy, dz, ix, iy, iz, col, maxSteps);
					} else if (dx == 0.f && dy == 0.f && dz == 0.f) {
						// no axis to cross: every one of the maxSteps iterations stays in the (empty) start cell with
						// t = 0 and the zero normal (:432-446 take no branch), then hit = false, hitPos = v0 (:503-506)
						write_hit(out, mine, 0u, v0x, v0y, v0z, ix, iy, iz, 0, 0, 0);
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
						// hasX = invX != 0 (:417): a NaN component counts as present, like there
						r.bX = r.invX != 0.f ? 0.f : kInf;
						r.bY = r.invY != 0.f ? 0.f : kInf;
						r.bZ = r.invZ != 0.f ? 0.f : kInf;
						r.sign = (px ? 1u : 0u) | (py ? 2u : 0u) | (pz ? 4u : 0u);
						r.ix = ix;
						r.iy = iy;
						r.iz = iz;
						r.col = col;
						r.left = maxSteps;
						r.ray = mine;
					}
				}
			}
			// ---- a few DDA steps between idle checks (:413-497) ----
#pragma unroll 1
			for (int k = 0; k < kStepsPerCheck; k++) {
				if (r.left != 0) {
					// the nearest boundary: X unless Y is strictly nearer, then Z if strictly nearer still (:432-446)
					const float tx = __fmaf_rn(r.fx, r.invX, r.bX), ty = __fmaf_rn(r.fy, r.invY, r.bY),
					            tz = __fmaf_rn(r.fz, r.invZ, r.bZ);
					const bool useY = ty < tx;
					float t = useY ? ty : tx;
					const bool isZ = tz < t;
					t = isZ ? tz : t;
					const bool isX = !useY && !isZ, isY = useY && !isZ;
					const float nfx = __fsub_rn(r.fx, __fmul_rn(r.ax, t));
					const float nfy = __fsub_rn(r.fy, __fmul_rn(r.ay, t));
					const float nfz = __fsub_rn(r.fz, __fmul_rn(r.az, t));
					r.fx = isX ? 1.f : nfx;
					r.fy = isY ? 1.f : nfy;
					r.fz = isZ ? 1.f : nfz;
					const int sx = (int)((r.sign << 1) & 2u) - 1, sy = (int)(r.sign & 2u) - 1, sz = (int)((r.sign >> 1) & 2u) - 1;
					r.ix += isX ? sx : 0;
					r.iy += isY ? sy : 0;
					r.iz += isZ ? sz : 0;
					r.left--;
					if (!isZ)
						r.col = solid[ray_index(r.ix, r.iy)];
					// IsSolidWrapped (GameMap.h:91-98) as one bit: bit z of the column, 0 above the map, 1 below it
					uint32_t bit = (uint32_t)(r.col >> (r.iz & 63));
					bit = (uint32_t)r.iz < 64u ? bit : ~((uint32_t)r.iz >> 31);
					const bool hit = (bit & 1u) != 0u;
					if (hit || r.left == 0) {
						// normal = iv - next = minus the step just taken
						const int nx = isX ? -sx : 0, ny = isY ? -sy : 0, nz = isZ ? -sz : 0;
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
						r.left = 0;
					}
				}
			}
		}
	}
} // namespace

namespace osb {
	// the solid map as the ray kernel reads it; brought up to date (one 2 MiB pass, no allocation) when the resident map
	// changed since the last ray call of this context
	int ray_map(osb200_ctx *c, const uint64_t **out) { // declared in ctx.h: k_castray (ambient.cu) reads the same copy
		if (c->solidXYVersion != c->mapVersion) {
			k_ray_map_xy<<<OSB200_MAP_W * OSB200_MAP_H / 256, 256, 0, c->stream>>>(c->d_solid, c->d_solidXY);
			c->launches++;
			OSB_CUDA(cudaGetLastError());
			c->solidXYVersion = c->mapVersion;
		}
		*out = c->d_solidXY;
		return 0;
	}
} // namespace osb
using osb::ray_map;

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
	const uint64_t *solidMap = nullptr;
	if (int rc = ray_map(c, &solidMap))
		return rc;
	size_t blocks = std::min<size_t>((size_t)c->smCount * kCtasPerSm, (n + kThreads - 1) / kThreads);
	{
		osb::KLaunch kl(c, osb::K_CASTRAY2);
		k_castray2<<<(unsigned)blocks, kThreads, 0, c->stream>>>(solidMap, d_v0, d_dir, maxSteps,
		                                                       reinterpret_cast<uint32_t *>(d_out), n, c->d_rayCursor);
	}
	OSB_CUDA(cudaGetLastError());
	return 0;
}

// Host-buffer entry point.  The batch is cut into chunks that flow through three stages on three streams -- H2D of
// chunk k+1, the kernel of chunk k, D2H of chunk k-1 -- over a ring of kPipeSlots device slots, so the call runs at the
// speed of the slower PCIe direction (40 B/ray out) instead of copy + kernel + copy in series.
//   pinned caller buffers    (cudaHostAlloc / cudaHostRegister memory): chunks of kPipeRays rays are copied straight
//                            between the caller's arrays and the device slots, every copy truly asynchronous;
//   pageable caller buffers  (what the reference's callers hold: std::vector, the stack): a cudaMemcpyAsync from
//                            pageable memory is staged by the driver on the calling thread at a few GB/s and blocks
//                            it meanwhile.  Here a few host threads move the data through a pinned ring instead: every
//                            thread copies its stripe of chunk k+1 into the ring, the last one to finish enqueues that
//                            chunk's H2D + kernel + D2H, and every thread then copies its stripe of chunk k's results
//                            out of the ring -- host copies, PCIe in, kernel and PCIe out all overlap.
namespace {
	bool is_pinned(const void *p) {
		cudaPointerAttributes a;
		if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
			cudaGetLastError(); // pageable memory: older runtimes report an error instead of "unregistered"
			return false;
		}
		return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
	}
	int copy_threads() {
		static const int v = [] {
			const char *e = getenv("OSB200_COPY_THREADS");
			int n = e ? atoi(e) : 0;
			if (n <= 0) {
				const unsigned hw = std::thread::hardware_concurrency();
				n = hw >= 16 ? 6 : (hw >= 8 ? 4 : (hw >= 4 ? 2 : 1));
			}
			return n > 16 ? 16 : n;
		}();
		return v;
	}

	int castray2_pinned(osb200_ctx *c, osb::RayPipe *P, const float *v0, const float *dir, int maxSteps, osb200_rayhit *out,
	                    size_t n) {
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

	int castray2_staged(osb200_ctx *c, osb::RayPipe *P, const float *v0, const float *dir, int maxSteps, osb200_rayhit *out,
	                    size_t n) {
		const size_t chunk = osb::kStagedRays;
		const size_t inPad = (chunk * 12 + 255) & ~(size_t)255, outPad = (chunk * sizeof(osb200_rayhit) + 255) & ~(size_t)255;
		const size_t slotBytes = 2 * inPad + outPad;
		if (c->ensure_scratch(slotBytes * osb::kPipeSlots))
			return -2;
		if (P->h_ring_bytes < slotBytes * osb::kPipeSlots) {
			if (P->h_ring)
				cudaFreeHost(P->h_ring);
			P->h_ring = nullptr;
			P->h_ring_bytes = 0;
			OSB_CUDA(cudaHostAlloc(&P->h_ring, slotBytes * osb::kPipeSlots, cudaHostAllocDefault));
			P->h_ring_bytes = slotBytes * osb::kPipeSlots;
		}
		char *dBase = reinterpret_cast<char *>(c->d_scratch);
		char *hBase = reinterpret_cast<char *>(P->h_ring);
		const size_t nChunks = (n + chunk - 1) / chunk;
		const int T = (int)std::min<size_t>((size_t)copy_threads(), std::max<size_t>(1, n >> 14));
		OSB_CUDA(cudaEventRecord(P->computed[0], c->stream));
		OSB_CUDA(cudaStreamWaitEvent(P->in, P->computed[0], 0));

		std::vector<std::atomic<int>> arrived(nChunks);
		for (auto &a : arrived)
			a.store(0, std::memory_order_relaxed);
		std::atomic<size_t> launched{0}; // chunks whose GPU work has been enqueued
		std::atomic<int> failed{0};
		std::string failText;
		std::mutex failLock;

		// enqueue chunk k: ring -> device, kernel, device -> ring (called by exactly one thread per chunk, in chunk order)
		auto launch = [&](size_t k) -> int {
			const int s = (int)(k % osb::kPipeSlots);
			const size_t m = std::min(chunk, n - k * chunk);
			char *d = dBase + slotBytes * s, *h = hBase + slotBytes * s;
			if (k >= (size_t)osb::kPipeSlots)
				OSB_CUDA(cudaStreamWaitEvent(P->in, P->downloaded[s], 0));
			OSB_CUDA(cudaMemcpyAsync(d, h, m * 12, cudaMemcpyHostToDevice, P->in));
			OSB_CUDA(cudaMemcpyAsync(d + inPad, h + inPad, m * 12, cudaMemcpyHostToDevice, P->in));
			OSB_CUDA(cudaEventRecord(P->uploaded[s], P->in));
			OSB_CUDA(cudaStreamWaitEvent(c->stream, P->uploaded[s], 0));
			int rc = osb200_castray2_device(c, reinterpret_cast<float *>(d), reinterpret_cast<float *>(d + inPad), maxSteps,
			                                reinterpret_cast<osb200_rayhit *>(d + 2 * inPad), m);
			if (rc)
				return rc;
			OSB_CUDA(cudaEventRecord(P->computed[s], c->stream));
			OSB_CUDA(cudaStreamWaitEvent(P->out, P->computed[s], 0));
			OSB_CUDA(cudaMemcpyAsync(h + 2 * inPad, d + 2 * inPad, m * sizeof(osb200_rayhit), cudaMemcpyDeviceToHost, P->out));
			OSB_CUDA(cudaEventRecord(P->downloaded[s], P->out));
			return 0;
		};
		auto fail = [&](int rc) {
			std::lock_guard<std::mutex> g(failLock);
			if (!failed.load()) {
				failText = osb200_last_error();
				failed.store(rc ? rc : -2);
			}
		};
		// bytes [lo, hi) of a chunk-sized copy that thread t of T moves (64-byte granules)
		auto stripe = [&](size_t bytes, int t, size_t &lo, size_t &hi) {
			const size_t g = (bytes + 63) / 64;
			lo = std::min(bytes, (g * (size_t)t / (size_t)T) * 64);
			hi = std::min(bytes, (g * (size_t)(t + 1) / (size_t)T) * 64);
		};
		auto worker = [&](int t) {
			if (cudaSetDevice(c->device) != cudaSuccess) {
				fail(-2);
				return;
			}
			auto copy_out = [&](size_t k) -> bool {
				while (launched.load(std::memory_order_acquire) <= k) { // its events must have been recorded first
					if (failed.load())
						return false;
					std::this_thread::yield();
				}
				const int s = (int)(k % osb::kPipeSlots);
				if (cudaEventSynchronize(P->downloaded[s]) != cudaSuccess) {
					osb::set_error("osb200_castray2: waiting for chunk %zu failed: %s", k, cudaGetErrorString(cudaGetLastError()));
					fail(-2);
					return false;
				}
				const size_t m = std::min(chunk, n - k * chunk);
				size_t lo, hi;
				stripe(m * sizeof(osb200_rayhit), t, lo, hi);
				memcpy(reinterpret_cast<char *>(out + k * chunk) + lo, hBase + slotBytes * s + 2 * inPad + lo, hi - lo);
				return true;
			};
			for (size_t k = 0; k < nChunks; k++) {
				if (failed.load())
					return;
				// the ring slot is free: this thread copied chunk k - kPipeSlots out of it in an earlier iteration, and
				// the chunk's H2D had finished long before its results came back
				const int s = (int)(k % osb::kPipeSlots);
				const size_t m = std::min(chunk, n - k * chunk);
				size_t lo, hi;
				stripe(m * 12, t, lo, hi);
				char *h = hBase + slotBytes * s;
				memcpy(h + lo, reinterpret_cast<const char *>(v0 + k * chunk * 3) + lo, hi - lo);
				memcpy(h + inPad + lo, reinterpret_cast<const char *>(dir + k * chunk * 3) + lo, hi - lo);
				if (arrived[k].fetch_add(1, std::memory_order_acq_rel) == T - 1) { // the last stripe of the chunk is in
					if (int rc = launch(k)) {
						fail(rc);
						return;
					}
					launched.store(k + 1, std::memory_order_release);
				}
				if (k >= 1 && !copy_out(k - 1))
					return;
			}
			copy_out(nChunks - 1);
		};
		std::vector<std::thread> pool;
		for (int t = 1; t < T; t++)
			pool.emplace_back(worker, t);
		worker(0);
		for (auto &th : pool)
			th.join();
		if (int rc = failed.load()) {
			osb::set_error("%s", failText.c_str());
			cudaStreamSynchronize(P->out);
			return rc;
		}
		OSB_CUDA(cudaStreamWaitEvent(c->stream, P->downloaded[(nChunks - 1) % osb::kPipeSlots], 0));
		return 0;
	}
} // namespace

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
	// a small batch is one chunk either way: the driver's own staging is as good and needs no threads
	const bool direct = n <= ((size_t)1 << 15) || (is_pinned(v0) && is_pinned(dir) && is_pinned(out));
	return direct ? castray2_pinned(c, c->rayPipe, v0, dir, maxSteps, out, n)
	              : castray2_staged(c, c->rayPipe, v0, dir, maxSteps, out, n);
}

// ---- CastRay2 without any allocation -------------------------------------------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// osb200_castray2 stages the caller's host buffers in device scratch that it allocates on demand (ensure_scratch, the ray
// cursor, the pipeline's streams and events).  This variant runs the same kernel with that scratch in a file-scope device
// array of fixed size: the batch cursor, the two input arrays and the result records of up to
// OSB200_CASTRAY2_STATIC_MAX_RAYS rays.  Every GPU that loads the library holds the array (4 MiB + 256 B) until exit.
//   * the array is shared by every call on a device and keeps the last call's contents: the cursor is cleared per call,
//     the input slots are overwritten by this call's copies, and the kernel writes every field of records 0 .. n-1;
//   * calls that overlap on one GPU (two contexts / streams / host threads) race on it -- serialise them.
namespace {
	constexpr size_t kStaticMaxRays = OSB200_CASTRAY2_STATIC_MAX_RAYS;
	constexpr size_t kStaticCursorBytes = 256;
	constexpr size_t kStaticInBytes = kStaticMaxRays * 3 * sizeof(float);       // per input array; a multiple of 256
	constexpr size_t kStaticOutBytes = kStaticMaxRays * sizeof(osb200_rayhit);  // a multiple of 256
	static_assert(kStaticInBytes % 256 == 0 && kStaticOutBytes % 256 == 0, "slots of the static scratch stay 256-byte aligned");
	static __device__ __align__(256) unsigned char g_rayScratch[kStaticCursorBytes + 2 * kStaticInBytes + kStaticOutBytes];

	// In host code `g_rayScratch` is a host-side stand-in, not a device address: the address a kernel may use comes from
	// cudaGetSymbolAddress, asked once per device with that device current.
	constexpr int kMaxDevices = 64;
	std::mutex g_rayScratchLock;
	unsigned char *g_rayScratchAddr[kMaxDevices] = {};

	int ray_scratch_of(int device, unsigned char **out) {
		if (device < 0 || device >= kMaxDevices) {
			osb::set_error("osb200_castray2_static: device %d beyond the %d devices this entry point keeps an address for", device,
			               kMaxDevices);
			return -1;
		}
		std::lock_guard<std::mutex> g(g_rayScratchLock);
		if (!g_rayScratchAddr[device]) {
			OSB_CUDA(cudaSetDevice(device));
			void *p = nullptr;
			OSB_CUDA(cudaGetSymbolAddress(&p, g_rayScratch));
			if (!p) {
				osb::set_error("osb200_castray2_static: cudaGetSymbolAddress returned a null address on device %d", device);
				return -2;
			}
			g_rayScratchAddr[device] = static_cast<unsigned char *>(p);
		}
		*out = g_rayScratchAddr[device];
		return 0;
	}
} // namespace

int osb200_castray2_static(osb200_ctx *c, const float *v0, const float *dir, int maxSteps, osb200_rayhit *out, size_t n) {
	if (!c || !c->mapLoaded) {
		osb::set_error("osb200_castray2_static: no context or no map uploaded");
		return -1;
	}
	if (n == 0)
		return 0;
	if (!v0 || !dir || !out) {
		osb::set_error("osb200_castray2_static: null buffer");
		return -1;
	}
	if (n > kStaticMaxRays) {
		osb::set_error("osb200_castray2_static: %zu rays exceed OSB200_CASTRAY2_STATIC_MAX_RAYS = %zu, the capacity of the static "
		               "device scratch; split the batch or use osb200_castray2 (which allocates)",
		               n, kStaticMaxRays);
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	unsigned char *base = nullptr;
	if (int rc = ray_scratch_of(c->device, &base))
		return rc;
	auto *cursor = reinterpret_cast<unsigned long long *>(base);
	auto *d_v0 = reinterpret_cast<float *>(base + kStaticCursorBytes);
	auto *d_dir = reinterpret_cast<float *>(base + kStaticCursorBytes + kStaticInBytes);
	auto *d_out = reinterpret_cast<uint32_t *>(base + kStaticCursorBytes + 2 * kStaticInBytes);
	cudaStream_t st = c->stream;
	OSB_CUDA(cudaMemsetAsync(cursor, 0, sizeof(unsigned long long), st)); // the one thing that must start at zero
	OSB_CUDA(cudaMemcpyAsync(d_v0, v0, n * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
	OSB_CUDA(cudaMemcpyAsync(d_dir, dir, n * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
	const size_t blocks = std::min<size_t>((size_t)c->smCount * kCtasPerSm, (n + kThreads - 1) / kThreads);
	const uint64_t *solidMap = nullptr;
	if (int rc = ray_map(c, &solidMap))
		return rc;
	k_castray2<<<(unsigned)blocks, kThreads, 0, st>>>(solidMap, d_v0, d_dir, maxSteps, d_out, n, cursor);
	c->launches++;
	OSB_CUDA(cudaGetLastError());
	OSB_CUDA(cudaMemcpyAsync(out, d_out, n * sizeof(osb200_rayhit), cudaMemcpyDeviceToHost, st));
	OSB_CUDA(cudaStreamSynchronize(st));
	return 0;
}

} // extern "C"
