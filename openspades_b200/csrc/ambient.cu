// GameMap::CastRay batch and the CPU half of the reference's ray-traced ambient-occlusion volume on the GPU:
//   GameMap::CastRay                         Sources/Client/GameMap.cpp:208-359        (cast_ray, k_castray)
//   GLAmbientShadowRenderer::Evaluate        Sources/Draw/GLAmbientShadowRenderer.cpp:130-169
//   GLAmbientShadowRenderer::GameMapChanged / Invalidate                               :171-237  (host, ambient_mark_dirty)
//   GLAmbientShadowRenderer::UpdateDirtyChunks (selection policy)                      :266-332  (host, select)
//   GLAmbientShadowRenderer::UpdateChunk     (window, compensation, 2 x 3 in-place blurs, copy-out)     :334-507  (k_ao_chunk)
//
// Design: one CTA per dirty chunk.  Everything a chunk's rays can touch is a 52 x 52 patch of column masks
// (window = chunk + 2 padding, rays walk at most 16 cell walls, contact test +-1): 21.6 KB, staged once into shared
// memory, so the 16 x <=16 DDA steps of every voxel never leave the SM.  16 lanes share one voxel (one ray each; the
// brightness terms are then added in ray order by shuffles, which keeps the reference's float summation order);
// the blur passes run one thread per line of the window because each tap reads its already-blurred predecessor.
// All float arithmetic is separate IEEE ops (--fmad=false); the one double multiply of Evaluate is kept in double.
#include "ctx.h"
#include "numerics.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <random>
#include <vector>

namespace osb {

	constexpr int kAoChunk = 16, kAoPad = 2, kAoWin = kAoChunk + 2 * kAoPad; // GLAmbientShadowRenderer.h:42-43, .cpp:344
	constexpr int kAoRays = 16, kAoRayLen = 16;                               // .h:41,44
	constexpr int kAoReach = kAoPad + kAoRayLen;                              // patch margin around the chunk: 18 columns
	constexpr int kAoPatch = kAoChunk + 2 * kAoReach;                         // 52
	constexpr int kAoChunks = 32 * 32 * 4;
	constexpr int kAoThreads = 512;

	struct AoTask {
		int cx, cy, cz;
		int lo[3], hi[3]; // dirty box inside the chunk, inclusive
	};

	struct AmbientState {
		float rays[kAoRays * 3];
		bool raysSet = false;
		float *d_rays = nullptr;
		// texture-layout volume: float2[65][512][512], slice z+1 holds map depth z, slice 0 stays (1,1)
		// (GLAmbientShadowRenderer.cpp:93-105,255-262: TexImage3D w x h x (d+1), chunks uploaded at cz*16 + 1)
		float2 *d_volume = nullptr;
		AoTask *d_tasks = nullptr, *h_tasks = nullptr; // pinned staging
		float *d_chunkOut = nullptr;                   // read_chunk gather target
		// chunk state exactly as the reference keeps it: {dirty, minX, minY, minZ, maxX, maxY, maxZ}, index (cx + cy*32)*4 + cz
		std::vector<int> state;
	};

	// Column mask fetch for the two callers of cast_ray: the tiled resident bitmap, or a staged patch in shared memory.
	struct GlobalColumns {
		const uint64_t *solid;
		__device__ __forceinline__ uint64_t operator()(int x, int y) const { return solid[tiled_index(x & 511, y & 511)]; }
	};
	struct LinearColumns { // the [x][y] copy of the map (ctx.h: ray_map): a cheaper address for the batch ray kernel
		const uint64_t *solidXY;
		__device__ __forceinline__ uint64_t operator()(int x, int y) const { return solidXY[((uint32_t)(x & 511) << 9) | (uint32_t)(y & 511)]; }
	};
	struct PatchColumns {
		const uint64_t *patch; // [kAoPatch][kAoPatch], origin (px0, py0) in world columns (unwrapped)
		int px0, py0;
		__device__ __forceinline__ uint64_t operator()(int x, int y) const { return patch[(x - px0) * kAoPatch + (y - py0)]; }
	};

	// GameMap::CastRay, the compiled "faster version" (GameMap.cpp:208-340).  The walk takes at most (long)length
	// cell walls; which wall comes next is a sign test on three integer cross-product accumulators.  Above the map
	// (a.z < 0) the reference neither tests voxels nor refreshes the cached column mask -- reproduced as is.
	template <class Columns>
	__device__ __forceinline__ bool cast_ray(const Columns &col, float v0x, float v0y, float v0z, float dx, float dy,
	                                         float dz, float length, int &hx, int &hy, int &hz) {
		const float v0[3] = {v0x, v0y, v0z};
		const float v1[3] = {__fadd_rn(v0x, __fmul_rn(dx, length)), __fadd_rn(v0y, __fmul_rn(dy, length)),
		                     __fadd_rn(v0z, __fmul_rn(dz, length))};
		int a[3], c[3], d[3];
		float f[3], g[3];
		long long cnt = 0;
#pragma unroll
		for (int k = 0; k < 3; k++) {
			a[k] = osbn::x86_f2i(floorf(v0[k]));
			c[k] = osbn::x86_f2i(floorf(v1[k]));
			if (c[k] < a[k]) {
				d[k] = -1;
				f[k] = __fsub_rn(v0[k], (float)a[k]);
				g[k] = __fmul_rn(__fsub_rn(v0[k], v1[k]), 1024.f);
				cnt += (long long)a[k] - c[k];
			} else if (c[k] != a[k]) {
				d[k] = 1;
				f[k] = __fsub_rn((float)(a[k] + 1), v0[k]);
				g[k] = __fmul_rn(__fsub_rn(v1[k], v0[k]), 1024.f);
				cnt += (long long)c[k] - a[k];
			} else {
				d[k] = 0;
				f[k] = g[k] = 0.f;
			}
		}
		int px = osbn::x86_f2i(floorf(__fsub_rn(__fmul_rn(f[0], g[2]), __fmul_rn(f[2], g[0]))));
		int py = osbn::x86_f2i(floorf(__fsub_rn(__fmul_rn(f[1], g[2]), __fmul_rn(f[2], g[1]))));
		int pz = osbn::x86_f2i(floorf(__fsub_rn(__fmul_rn(f[1], g[0]), __fmul_rn(f[0], g[1]))));
		const int ix = osbn::x86_f2i(floorf(g[0])), iy = osbn::x86_f2i(floorf(g[1])), iz = osbn::x86_f2i(floorf(g[2]));
		// (long)length: cvttss2si r64
		const long long lim = (length >= -9223372036854775808.f && length < 9223372036854775808.f)
		                        ? (long long)length
		                        : (long long)0x8000000000000000ull;
		if (cnt > lim)
			cnt = lim;
		int ax = a[0], ay = a[1], az = a[2];
		uint64_t mask = col(ax, ay);
		if (az < 0 && d[2] < 0)
			return false;
		if (az >= 64) {
			hx = ax, hy = ay, hz = az;
			return true;
		}
		while (cnt > 0) {
			const bool above = az < 0;
			if (((px | py) >= 0) && az != c[2]) {
				az += d[2];
				px -= ix;
				py -= iy;
				if (!above) {
					if (az < 0 && d[2] < 0)
						return false;
					if (az >= 64) {
						hx = ax, hy = ay, hz = az;
						return true;
					}
				}
			} else {
				if (pz >= 0 && ax != c[0]) {
					ax += d[0];
					px += iz;
					pz -= iy;
				} else {
					ay += d[1];
					py += iz;
					pz += ix;
				}
				if (!above)
					mask = col(ax, ay);
			}
			if (!above && ((mask >> az) & 1ull)) {
				hx = ax, hy = ay, hz = az;
				return true;
			}
			cnt--;
		}
		return false;
	}

	// The same walk as a resumable state, for the batch kernel: lanes whose ray has ended pick up new rays while the
	// others keep walking (rays of one batch end after anything between 0 and ~2 * length cell walls).
	struct RayWalk {
		int ax, ay, az, px, py, pz, ix, iy, iz, dx, dy, dz, cx, cz;
		long long cnt;
		uint64_t mask;
	};
	enum { kRayWalking = 0, kRayHit = 1, kRayMiss = 2 };
	template <class Columns>
	__device__ __forceinline__ int ray_init(const Columns &col, const float v0[3], const float dir[3], float length, RayWalk &w) {
		const float v1[3] = {__fadd_rn(v0[0], __fmul_rn(dir[0], length)), __fadd_rn(v0[1], __fmul_rn(dir[1], length)),
		                     __fadd_rn(v0[2], __fmul_rn(dir[2], length))};
		int a[3], c[3], d[3];
		float f[3], g[3];
		long long cnt = 0;
#pragma unroll
		for (int k = 0; k < 3; k++) {
			a[k] = osbn::x86_f2i(floorf(v0[k]));
			c[k] = osbn::x86_f2i(floorf(v1[k]));
			if (c[k] < a[k]) {
				d[k] = -1;
				f[k] = __fsub_rn(v0[k], (float)a[k]);
				g[k] = __fmul_rn(__fsub_rn(v0[k], v1[k]), 1024.f);
				cnt += (long long)a[k] - c[k];
			} else if (c[k] != a[k]) {
				d[k] = 1;
				f[k] = __fsub_rn((float)(a[k] + 1), v0[k]);
				g[k] = __fmul_rn(__fsub_rn(v1[k], v0[k]), 1024.f);
				cnt += (long long)c[k] - a[k];
			} else {
				d[k] = 0;
				f[k] = g[k] = 0.f;
			}
		}
		w.px = osbn::x86_f2i(floorf(__fsub_rn(__fmul_rn(f[0], g[2]), __fmul_rn(f[2], g[0]))));
		w.py = osbn::x86_f2i(floorf(__fsub_rn(__fmul_rn(f[1], g[2]), __fmul_rn(f[2], g[1]))));
		w.pz = osbn::x86_f2i(floorf(__fsub_rn(__fmul_rn(f[1], g[0]), __fmul_rn(f[0], g[1]))));
		w.ix = osbn::x86_f2i(floorf(g[0]));
		w.iy = osbn::x86_f2i(floorf(g[1]));
		w.iz = osbn::x86_f2i(floorf(g[2]));
		const long long lim = (length >= -9223372036854775808.f && length < 9223372036854775808.f)
		                        ? (long long)length
		                        : (long long)0x8000000000000000ull;
		w.cnt = cnt > lim ? lim : cnt;
		w.ax = a[0], w.ay = a[1], w.az = a[2];
		w.dx = d[0], w.dy = d[1], w.dz = d[2];
		w.cx = c[0], w.cz = c[2];
		w.mask = col(w.ax, w.ay);
		if (w.az < 0 && w.dz < 0)
			return kRayMiss;
		if (w.az >= 64)
			return kRayHit;
		return w.cnt > 0 ? kRayWalking : kRayMiss;
	}
	template <class Columns> __device__ __forceinline__ int ray_step(const Columns &col, RayWalk &w) {
		const bool above = w.az < 0;
		if (((w.px | w.py) >= 0) && w.az != w.cz) {
			w.az += w.dz;
			w.px -= w.ix;
			w.py -= w.iy;
			if (!above) {
				if (w.az < 0 && w.dz < 0)
					return kRayMiss;
				if (w.az >= 64)
					return kRayHit;
			}
		} else {
			if (w.pz >= 0 && w.ax != w.cx) {
				w.ax += w.dx;
				w.px += w.iz;
				w.pz -= w.iy;
			} else {
				w.ay += w.dy;
				w.py += w.iz;
				w.pz += w.ix;
			}
			if (!above)
				w.mask = col(w.ax, w.ay);
		}
		if (!above && ((w.mask >> w.az) & 1ull))
			return kRayHit;
		return --w.cnt > 0 ? kRayWalking : kRayMiss;
	}

	// Batch kernel: persistent warps over a batch-wide cursor; when half of a warp's lanes have no ray, the idle lanes
	// fetch the next rays (first version: one ray per thread to completion, 9.9 of 32 lanes busy on average).
#ifndef OSB_RAY_BURST
#define OSB_RAY_BURST 8
#endif
#ifndef OSB_RAY_REFILL
#define OSB_RAY_REFILL 24 // refill when at most this many lanes are walking (8 / 16 / 24 / 28 measured: 5.34 / 4.11 / 3.71 / 3.89 ms)
#endif
	constexpr int kRayBurst = OSB_RAY_BURST;
	__global__ void __launch_bounds__(256) k_castray(const uint64_t *__restrict__ solid, const float *__restrict__ v0,
	                                                const float *__restrict__ dir, float length, int4 *__restrict__ out,
	                                                size_t n, unsigned long long *__restrict__ cursor) {
		const LinearColumns col{solid}; // `solid` is the [x][y] copy here
		const unsigned lane = threadIdx.x & 31;
		RayWalk w;
		size_t idx = 0;
		bool active = false, more = true;
		for (;;) {
			unsigned act = __ballot_sync(0xffffffffu, active);
			if (more && __popc(act) <= OSB_RAY_REFILL) {
				const unsigned idle = ~act;
				const int need = __popc(idle);
				unsigned long long base = 0;
				if (lane == 0)
					base = atomicAdd(cursor, (unsigned long long)need);
				base = __shfl_sync(0xffffffffu, base, 0);
				more = base + (unsigned long long)need < n;
				if (!active) {
					const size_t i = base + __popc(idle & ((1u << lane) - 1u));
					if (i < n) {
						const float o[3] = {v0[i * 3], v0[i * 3 + 1], v0[i * 3 + 2]};
						const float d[3] = {dir[i * 3], dir[i * 3 + 1], dir[i * 3 + 2]};
						const int st = ray_init(col, o, d, length, w);
						if (st == kRayWalking) {
							active = true;
							idx = i;
						} else {
							out[i] = st == kRayHit ? make_int4(1, w.ax, w.ay, w.az) : make_int4(0, 0, 0, 0);
						}
					}
				}
				act = __ballot_sync(0xffffffffu, active);
			}
			if (!act) {
				if (!more)
					break;
				continue;
			}
#pragma unroll 1
			for (int s = 0; s < kRayBurst; s++) {
				if (active) {
					const int st = ray_step(col, w);
					if (st != kRayWalking) {
						out[idx] = st == kRayHit ? make_int4(1, w.ax, w.ay, w.az) : make_int4(0, 0, 0, 0);
						active = false;
					}
				}
			}
		}
	}

	// One ray of Evaluate (GLAmbientShadowRenderer.cpp:141-162): brightness of ray i from the centre of voxel (x, y, z).
	template <class Columns>
	__device__ __forceinline__ float ao_ray(const Columns &col, const float *__restrict__ rays, int i, int x, int y, int z) {
		const float px = __fadd_rn((float)x, 0.5f), py = __fadd_rn((float)y, 0.5f), pz = __fadd_rn((float)z, 0.5f);
		float dx = rays[i * 3], dy = rays[i * 3 + 1], dz = rays[i * 3 + 2];
		dx = (i & 1) ? -dx : dx;
		dy = (i & 2) ? -dy : dy;
		dz = (i & 4) ? -dz : dz;
		int hx, hy, hz;
		float brightness = 1.f;
		if (cast_ray(col, px, py, pz, dx, dy, dz, (float)kAoRayLen, hx, hy, hz)) {
			const float ex = __fsub_rn(__fadd_rn((float)hx, .5f), px), ey = __fsub_rn(__fadd_rn((float)hy, .5f), py),
			            ez = __fsub_rn(__fadd_rn((float)hz, .5f), pz);
			const float dist = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(ez, ez));
			// `dist * (1.0 / float(15 * 15))` is a double product rounded to float on assignment
			brightness = __double2float_rn(__dmul_rn((double)dist, 1.0 / 225.0));
			brightness = brightness > 1.f ? 1.f : brightness;
		}
		return brightness;
	}

	// 16 lanes (one half-warp) evaluate one voxel: lane r casts ray r, then every lane adds the 16 terms in ray order.
	template <class Columns>
	__device__ __forceinline__ float ao_evaluate16(const Columns &col, const float *__restrict__ rays, int x, int y, int z,
	                                               bool active) {
		const int r = threadIdx.x & 15;
		const float b = active ? ao_ray(col, rays, r, x, y, z) : 0.f;
		float sum = 0.f;
		const int base = threadIdx.x & 16; // lane of ray 0 inside the warp
#pragma unroll
		for (int i = 0; i < kAoRays; i++)
			sum = __fadd_rn(sum, __shfl_sync(0xffffffffu, b, base + i));
		const float s = __fmul_rn(sum, 2.f / (float)kAoRays);
		return s < 1.f ? s : 1.f;
	}

	__device__ __forceinline__ bool solid_wrapped(const PatchColumns &col, int x, int y, int z) { // GameMap.h:91-98
		if (z < 0)
			return false;
		if (z >= 64)
			return true;
		return (col(x, y) >> z) & 1ull;
	}

	// shouldBlur (:457-459): blur with a neighbour that is air and is a contact voxel, or when this voxel is not one
	__device__ __forceinline__ bool should_blur(uint32_t self, uint32_t nb) { return ((nb & 1u) | ((~self | nb) & 2u)) == 2u; }

	struct AoSmem {
		uint64_t patch[kAoPatch * kAoPatch];
		float w[kAoWin][kAoWin][kAoWin];
		uint8_t flags[kAoWin][kAoWin][kAoWin];
		float rays[kAoRays * 3];
		uint64_t reach[kAoWin][kAoWin]; // OR of the column masks within ray reach (+-16 columns) of a window column
		union {
			uint64_t reachY[kAoPatch][kAoWin];            // first half of the separable OR (along y)
			uint16_t list[kAoWin * kAoWin * kAoWin];      // window voxels that need rays, compacted
		} u;
		int listCount;
	};

	// UpdateChunk for one task per CTA.
	__global__ void __launch_bounds__(kAoThreads) k_ao_chunk(const uint64_t *__restrict__ solid,
	                                                        const float *__restrict__ rays,
	                                                        const AoTask *__restrict__ tasks, float2 *__restrict__ volume) {
		extern __shared__ __align__(16) unsigned char s_raw[];
		AoSmem &S = *reinterpret_cast<AoSmem *>(s_raw);
		const AoTask T = tasks[blockIdx.x];
		const int ox = T.cx * kAoChunk - kAoPad, oy = T.cy * kAoChunk - kAoPad, oz = T.cz * kAoChunk - kAoPad;
		const PatchColumns col{S.patch, T.cx * kAoChunk - kAoReach, T.cy * kAoChunk - kAoReach};
		for (int i = threadIdx.x; i < kAoPatch * kAoPatch; i += kAoThreads) {
			const int x = col.px0 + i / kAoPatch, y = col.py0 + i % kAoPatch;
			S.patch[i] = solid[tiled_index(x & 511, y & 511)];
		}
		if (threadIdx.x < kAoRays * 3)
			S.rays[threadIdx.x] = rays[threadIdx.x];
		__syncthreads();

		// window = dirty box grown by the padding on both sides (:352-357), in window coordinates
		const int x0 = T.lo[0], y0 = T.lo[1], z0 = T.lo[2];
		const int x1 = T.hi[0] + 2 * kAoPad, y1 = T.hi[1] + 2 * kAoPad, z1 = T.hi[2] + 2 * kAoPad;
		const int nx = x1 - x0 + 1, ny = y1 - y0 + 1, nz = z1 - z0 + 1;
		const int nvox = nx * ny * nz;

		// ---- flags: bit 0 solid, bit 1 contact.  The reference's neighbour list names (x-1,y-1,z-1) twice and never
		// (x,y-1,z-1) (:383-409): 25 of the 26 shell voxels.
		for (int i = threadIdx.x; i < nvox; i += kAoThreads) {
			const int x = x0 + i % nx, y = y0 + (i / nx) % ny, z = z0 + i / (nx * ny);
			const int wx = x + ox, wy = y + oy, wz = z + oz;
			bool contact = false;
#pragma unroll
			for (int dx = -1; dx <= 1; dx++)
#pragma unroll
				for (int dy = -1; dy <= 1; dy++)
#pragma unroll
					for (int dz = -1; dz <= 1; dz++) {
						if ((dx == 0 && dy == 0 && dz == 0) || (dx == 0 && dy == -1 && dz == -1))
							continue;
						contact |= solid_wrapped(col, wx + dx, wy + dy, wz + dz);
					}
			S.flags[z][y][x] = (uint8_t)((solid_wrapped(col, wx, wy, wz) ? 1 : 0) | (contact ? 2 : 0));
		}
		__syncthreads();

		// ---- which voxels need rays at all.  A ray crosses at most 16 cell walls, so it stays within +-16 cells of its
		// voxel on every axis; when no solid voxel (and not the solid half-space z >= 64) lies in that box all 16 rays
		// miss and Evaluate is exactly min(16 * 2/16, 1) = 1, which the compensation maps to 1.  The per-column OR over
		// the 33 x 33 neighbourhood is separable: along y into reachY, then along x into reach.
		for (int i = threadIdx.x; i < kAoPatch * kAoWin; i += kAoThreads) {
			const int px = i / kAoWin, wy = i % kAoWin;
			uint64_t m = 0;
			for (int dy = 0; dy <= 2 * kAoRayLen; dy++)
				m |= S.patch[px * kAoPatch + wy + dy];
			S.u.reachY[px][wy] = m;
		}
		if (threadIdx.x == 0)
			S.listCount = 0;
		__syncthreads();
		for (int i = threadIdx.x; i < kAoWin * kAoWin; i += kAoThreads) {
			const int wx = i / kAoWin, wy = i % kAoWin;
			uint64_t m = 0;
			for (int dx = 0; dx <= 2 * kAoRayLen; dx++)
				m |= S.u.reachY[wx + dx][wy];
			S.reach[wx][wy] = m;
		}
		__syncthreads();
		for (int i = threadIdx.x; i < nvox; i += kAoThreads) {
			const int x = x0 + i % nx, y = y0 + (i / nx) % ny, z = z0 + i / (nx * ny);
			const int wz = z + oz;
			if (S.flags[z][y][x] & 1) {
				S.w[z][y][x] = 0.f; // solid: 0 before and after the compensation
				continue;
			}
			const int lo = max(wz - kAoRayLen, 0), hi = min(wz + kAoRayLen, 63);
			bool clear = wz + kAoRayLen <= 63;
			if (clear && hi >= lo) {
				const uint64_t zmask = (hi == 63 ? ~0ull : ((1ull << (hi + 1)) - 1ull)) & ~((1ull << lo) - 1ull);
				clear = (S.reach[x][y] & zmask) == 0;
			}
			if (clear)
				S.w[z][y][x] = 1.f;
			else
				S.u.list[atomicAdd(&S.listCount, 1)] = (uint16_t)((z * kAoWin + y) * kAoWin + x);
		}
		__syncthreads();

		// ---- Evaluate for the listed voxels (16 lanes each) + the under-shadowing compensation d *= d*d + 1 - d (:417-423)
		{
			const int group = threadIdx.x >> 4, groups = kAoThreads >> 4;
			const int n = S.listCount;
			const int rounds = (n + groups - 1) / groups;
			for (int k = 0; k < rounds; k++) {
				const int i = k * groups + group;
				const bool inside = i < n;
				const int v = inside ? S.u.list[i] : 0;
				const int x = v % kAoWin, y = (v / kAoWin) % kAoWin, z = v / (kAoWin * kAoWin);
				float d = ao_evaluate16(col, S.rays, x + ox, y + oy, z + oz, inside);
				if (inside && (threadIdx.x & 15) == 0) {
					d = __fmul_rn(d, __fsub_rn(__fadd_rn(__fmul_rn(d, d), 1.0f), d));
					S.w[z][y][x] = d;
				}
			}
		}
		__syncthreads();

		// ---- two rounds of in-place 3-tap blurs along x, y, z (:461-500).  A tap reads its predecessor after that one
		// was blurred, so a line is a serial chain: one thread per line, lines in parallel.
		for (int pass = 0; pass < 2; pass++) {
			for (int axis = 0; axis < 3; axis++) {
				// axis 0: lines over (z, y) stepping x;  1: over (z, x) stepping y;  2: over (y, x) stepping z
				const int nA = axis == 0 ? ny : nx, nB = axis == 2 ? ny : nz;
				const int len = axis == 0 ? nx : (axis == 1 ? ny : nz);
				for (int l = threadIdx.x; l < nA * nB; l += kAoThreads) {
					const int ia = l % nA, ib = l / nA;
					int x = x0, y = y0, z = z0, sx = 0, sy = 0, sz = 0;
					if (axis == 0) {
						y += ia, z += ib, sx = 1;
					} else if (axis == 1) {
						x += ia, z += ib, sy = 1;
					} else {
						x += ia, y += ib, sz = 1;
					}
					for (int t = 1; t < len - 1; t++) {
						const int cx = x + sx * t, cy = y + sy * t, cz = z + sz * t;
						const uint32_t self = S.flags[cz][cy][cx];
						if (self & 1u)
							continue;
						const bool m1 = should_blur(self, S.flags[cz - sz][cy - sy][cx - sx]);
						const bool m2 = should_blur(self, S.flags[cz + sz][cy + sy][cx + sx]);
						float v = S.w[cz][cy][cx];
						v = __fadd_rn(v, m1 ? S.w[cz - sz][cy - sy][cx - sx] : 0.0f);
						v = __fadd_rn(v, m2 ? S.w[cz + sz][cy + sy][cx + sx] : 0.0f);
						S.w[cz][cy][cx] = __fmul_rn(v, (m1 && m2) ? 1.0f / 3.0f : ((m1 || m2) ? 1.0f / 2.0f : 1.0f)); // divider[] :455
					}
				}
				__syncthreads();
			}
		}

		// ---- copy the dirty box out (:502-508), texture layout: slice = world z + 1, x / y inside the map
		const int bx = T.hi[0] - T.lo[0] + 1, by = T.hi[1] - T.lo[1] + 1, bz = T.hi[2] - T.lo[2] + 1;
		for (int i = threadIdx.x; i < bx * by * bz; i += kAoThreads) {
			const int x = T.lo[0] + i % bx, y = T.lo[1] + (i / bx) % by, z = T.lo[2] + i / (bx * by);
			const float ao = S.w[z + kAoPad][y + kAoPad][x + kAoPad];
			const float air = (S.flags[z + kAoPad][y + kAoPad][x + kAoPad] & 1) ? 0.f : 1.f;
			const size_t o = ((size_t)(T.cz * kAoChunk + z + 1) * 512 + (T.cy * kAoChunk + y)) * 512 + (T.cx * kAoChunk + x);
			volume[o] = make_float2(ao, air);
		}
	}

	// Evaluate over a box (parity hook): 16 lanes per voxel, column masks straight from the resident bitmap
	__global__ void __launch_bounds__(256) k_ao_evaluate_box(const uint64_t *__restrict__ solid, const float *__restrict__ rays,
	                                                        int x0, int y0, int z0, int nx, int ny, int nz,
	                                                        float *__restrict__ out) {
		const GlobalColumns col{solid};
		const size_t nvox = (size_t)nx * ny * nz;
		const size_t group = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
		const bool inside = group < nvox;
		const size_t i = inside ? group : 0;
		const int x = x0 + (int)(i % nx), y = y0 + (int)((i / nx) % ny), z = z0 + (int)(i / ((size_t)nx * ny));
		const float v = ao_evaluate16(col, rays, x, y, z, inside);
		if (inside && (threadIdx.x & 15) == 0)
			out[i] = v;
	}

	__global__ void k_fill_f2(float2 *p, size_t n, float2 v) {
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
			p[i] = v;
	}
	__global__ void k_ao_gather_chunk(const float2 *__restrict__ vol, int X0, int Y0, int Z0, float2 *__restrict__ out) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x; // [z][y][x]
		const int lx = i & 15, ly = (i >> 4) & 15, lz = i >> 8;
		out[i] = vol[((size_t)(Z0 + lz + 1) * 512 + (Y0 + ly)) * 512 + (X0 + lx)];
	}

	// ---- host side -------------------------------------------------------------------------------

	static void ambient_reset_state(AmbientState *A) {
		// a freshly constructed GLAmbientShadowRenderer: every chunk dirty over its whole 16^3 (.h:55-66)
		A->state.assign((size_t)kAoChunks * 7, 0);
		for (int i = 0; i < kAoChunks; i++) {
			int *s = &A->state[(size_t)i * 7];
			s[0] = 1;
			s[4] = s[5] = s[6] = kAoChunk - 1;
		}
	}

	static int ambient_init(osb200_ctx *c) {
		if (c->ambient)
			return 0;
		auto *A = new AmbientState();
		c->ambient = A;
		// rays: Normalize(random in [0,1)^3) + 0.01 (:53-60).  The reference draws them per instance from its global
		// RNG, so they are an input here (osb200_ambient_set_rays); these defaults follow the same recipe.
		std::mt19937 rng(0x5eed);
		std::uniform_real_distribution<float> U(0.f, 1.f);
		for (int i = 0; i < kAoRays; i++) {
			float v[3] = {U(rng), U(rng), U(rng)};
			float len = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
			if (len != 0.f)
				len = 1.f / len;
			for (int k = 0; k < 3; k++)
				A->rays[i * 3 + k] = v[k] * len + 0.01f;
		}
		const size_t nvol = (size_t)512 * 512 * 65;
		OSB_CUDA(cudaMalloc(&A->d_rays, sizeof(A->rays)));
		OSB_CUDA(cudaMalloc(&A->d_volume, nvol * sizeof(float2)));
		OSB_CUDA(cudaMalloc(&A->d_tasks, sizeof(AoTask) * kAoChunks));
		OSB_CUDA(cudaMallocHost(&A->h_tasks, sizeof(AoTask) * kAoChunks));
		OSB_CUDA(cudaMalloc(&A->d_chunkOut, sizeof(float2) * 4096));
		OSB_CUDA(cudaMemcpyAsync(A->d_rays, A->rays, sizeof(A->rays), cudaMemcpyHostToDevice, c->stream));
		k_fill_f2<<<c->smCount * 8, 256, 0, c->stream>>>(A->d_volume, nvol, make_float2(1.f, 1.f)); // (:69-72,100-105)
		c->launches++;
		OSB_CUDA(cudaFuncSetAttribute(k_ao_chunk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AoSmem)));
		ambient_reset_state(A);
		OSB_CUDA(cudaGetLastError());
		return 0;
	}

	void ambient_destroy(osb200_ctx *c) {
		AmbientState *A = c->ambient;
		if (!A)
			return;
		cudaFree(A->d_rays);
		cudaFree(A->d_volume);
		cudaFree(A->d_tasks);
		cudaFreeHost(A->h_tasks);
		cudaFree(A->d_chunkOut);
		delete A;
		c->ambient = nullptr;
	}

	// osb200_map_upload replaced the map: same as constructing a new GLAmbientShadowRenderer for it
	void ambient_map_replaced(osb200_ctx *c) {
		AmbientState *A = c->ambient;
		if (!A)
			return;
		ambient_reset_state(A);
		k_fill_f2<<<c->smCount * 8, 256, 0, c->stream>>>(A->d_volume, (size_t)512 * 512 * 65, make_float2(1.f, 1.f));
		c->launches++;
	}

	// GLAmbientShadowRenderer::GameMapChanged -> Invalidate(+-RayLength) (:171-237): host-only, a few chunk boxes
	void ambient_mark_dirty(osb200_ctx *c, int x, int y, int z) {
		AmbientState *A = c->ambient;
		if (!A)
			return; // not in use yet: the first use starts with every chunk dirty anyway
		int lo[3] = {x - kAoRayLen, y - kAoRayLen, z - kAoRayLen}, hi[3] = {x + kAoRayLen, y + kAoRayLen, z + kAoRayLen};
		lo[2] = std::max(lo[2], 0);
		hi[2] = std::min(hi[2], 63);
		if (lo[2] > hi[2])
			return;
		for (int cx = lo[0] >> 4; cx <= (hi[0] >> 4); cx++)
			for (int cy = lo[1] >> 4; cy <= (hi[1] >> 4); cy++)
				for (int cz = lo[2] >> 4; cz <= (hi[2] >> 4); cz++) {
					int *s = &A->state[(size_t)(((cx & 31) + (cy & 31) * 32) * 4 + cz) * 7];
					const int org[3] = {cx * kAoChunk, cy * kAoChunk, cz * kAoChunk};
					for (int k = 0; k < 3; k++) {
						const int mn = std::max(lo[k] - org[k], 0), mx = std::min(hi[k] - org[k], kAoChunk - 1);
						s[1 + k] = s[0] ? std::min(s[1 + k], mn) : mn;
						s[4 + k] = s[0] ? std::max(s[4 + k], mx) : mx;
					}
					s[0] = 1;
				}
	}
} // namespace osb

extern "C" {

int osb200_castray_device(osb200_ctx *c, const float *d_v0, const float *d_dir, float length, int32_t *d_out, size_t n) {
	if (!c || !c->mapLoaded) {
		osb::set_error("osb200_castray: no context or no map uploaded");
		return -1;
	}
	if (n == 0)
		return 0;
	if (!d_v0 || !d_dir || !d_out) {
		osb::set_error("osb200_castray: null buffer");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	const size_t blocks = std::min<size_t>((size_t)c->smCount * 8, (n + 255) / 256);
	if (!c->d_rayCursor)
		OSB_CUDA(cudaMalloc(&c->d_rayCursor, sizeof(unsigned long long)));
	OSB_CUDA(cudaMemsetAsync(c->d_rayCursor, 0, sizeof(unsigned long long), c->stream));
	const uint64_t *solidXY = nullptr;
	if (int rc = osb::ray_map(c, &solidXY))
		return rc;
	{
		osb::KLaunch kl(c, osb::K_CASTRAY);
		osb::k_castray<<<(unsigned)blocks, 256, 0, c->stream>>>(solidXY, d_v0, d_dir, length,
		                                                       reinterpret_cast<int4 *>(d_out), n, c->d_rayCursor);
	}
	OSB_CUDA(cudaGetLastError());
	return 0;
}

int osb200_castray(osb200_ctx *c, const float *v0, const float *dir, float length, int32_t *out, size_t n) {
	if (!c || !c->mapLoaded) {
		osb::set_error("osb200_castray: no context or no map uploaded");
		return -1;
	}
	if (n == 0)
		return 0;
	if (!v0 || !dir || !out) {
		osb::set_error("osb200_castray: null buffer");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	const size_t inB = n * 12, outB = n * 16;
	const size_t inPad = (inB + 255) & ~(size_t)255;
	if (c->ensure_scratch(2 * inPad + outB))
		return -2;
	char *p = reinterpret_cast<char *>(c->d_scratch);
	float *dv = reinterpret_cast<float *>(p), *dd = reinterpret_cast<float *>(p + inPad);
	int32_t *dout = reinterpret_cast<int32_t *>(p + 2 * inPad);
	OSB_CUDA(cudaMemcpyAsync(dv, v0, inB, cudaMemcpyHostToDevice, c->stream));
	OSB_CUDA(cudaMemcpyAsync(dd, dir, inB, cudaMemcpyHostToDevice, c->stream));
	int rc = osb200_castray_device(c, dv, dd, length, dout, n);
	if (rc)
		return rc;
	OSB_CUDA(cudaMemcpyAsync(out, dout, outB, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_ambient_set_rays(osb200_ctx *c, const float rays[48]) {
	if (!c || !rays) {
		osb::set_error("osb200_ambient_set_rays: null argument");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::ambient_init(c))
		return -2;
	osb::AmbientState *A = c->ambient;
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	std::memcpy(A->rays, rays, sizeof(A->rays));
	OSB_CUDA(cudaMemcpyAsync(A->d_rays, A->rays, sizeof(A->rays), cudaMemcpyHostToDevice, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_ambient_get_rays(osb200_ctx *c, float rays[48]) {
	if (!c || !rays) {
		osb::set_error("osb200_ambient_get_rays: null argument");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::ambient_init(c))
		return -2;
	std::memcpy(rays, c->ambient->rays, sizeof(c->ambient->rays));
	return 0;
}

int osb200_ambient_update(osb200_ctx *c, int maxChunks, const float *eye) {
	if (!c || !c->mapLoaded) {
		osb::set_error("osb200_ambient_update: no context or no map uploaded");
		return -1;
	}
	if (c->anyDirty) {
		osb::set_error("ambient update: edited columns are pending -- call osb200_map_flush first");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::ambient_init(c))
		return -2;
	osb::AmbientState *A = c->ambient;
	// ---- selection (UpdateDirtyChunks :266-332): dirty chunks within 6 chunks of the eye first, any dirty chunk only
	// when there is none; the reference then draws 8 of (at most 256) candidates at random, this takes the first
	// maxChunks in chunk-index order (maxChunks < 0: every dirty chunk)
	std::vector<int> pick;
	auto scan = [&](bool nearOnly) {
		int ex = 0, ey = 0, ez = 0;
		if (nearOnly) {
			ex = (int)eye[0] >> 4;
			ey = (int)eye[1] >> 4;
			ez = (int)eye[2] >> 4;
		}
		for (int i = 0; i < osb::kAoChunks; i++) {
			if (!A->state[(size_t)i * 7])
				continue;
			if (nearOnly) {
				const int cz = i & 3, cx = (i >> 2) & 31, cy = i >> 7;
				const int dx = (cx - ex) & 31, dy = (cy - ey) & 31, dz = cz - ez;
				if ((dx >= 6 && dx <= 32 - 6) || (dy >= 6 && dy <= 32 - 6) || dz >= 6 || dz <= -6)
					continue;
			}
			pick.push_back(i);
		}
	};
	if (eye && maxChunks >= 0)
		scan(true);
	if (pick.empty())
		scan(false);
	if (maxChunks >= 0 && (int)pick.size() > maxChunks)
		pick.resize(maxChunks);
	const int n = (int)pick.size();
	if (n == 0)
		return 0;
	OSB_CUDA(cudaStreamSynchronize(c->stream)); // h_tasks may still be in flight
	for (int k = 0; k < n; k++) {
		const int i = pick[k];
		int *s = &A->state[(size_t)i * 7];
		osb::AoTask &t = A->h_tasks[k];
		t.cz = i & 3;
		t.cx = (i >> 2) & 31;
		t.cy = i >> 7;
		for (int j = 0; j < 3; j++) {
			t.lo[j] = s[1 + j];
			t.hi[j] = s[4 + j];
		}
		s[0] = 0;
	}
	OSB_CUDA(cudaMemcpyAsync(A->d_tasks, A->h_tasks, sizeof(osb::AoTask) * n, cudaMemcpyHostToDevice, c->stream));
	{
		osb::KLaunch kl(c, osb::K_AMBIENT);
		osb::k_ao_chunk<<<n, osb::kAoThreads, sizeof(osb::AoSmem), c->stream>>>(c->d_solid, A->d_rays, A->d_tasks,
		                                                                         A->d_volume);
	}
	OSB_CUDA(cudaGetLastError());
	return n;
}

int osb200_ambient_chunk_info(osb200_ctx *c, int cx, int cy, int cz, int out[7]) {
	if (!c || !out || cx < 0 || cx >= 32 || cy < 0 || cy >= 32 || cz < 0 || cz >= 4) {
		osb::set_error("osb200_ambient_chunk_info: bad argument");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::ambient_init(c))
		return -2;
	std::memcpy(out, &c->ambient->state[(size_t)((cx + cy * 32) * 4 + cz) * 7], 7 * sizeof(int));
	return 0;
}

int osb200_ambient_read_chunk(osb200_ctx *c, int cx, int cy, int cz, float *out) {
	if (!c || !out || cx < 0 || cx >= 32 || cy < 0 || cy >= 32 || cz < 0 || cz >= 4) {
		osb::set_error("osb200_ambient_read_chunk: bad argument");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::ambient_init(c))
		return -2;
	osb::AmbientState *A = c->ambient;
	osb::k_ao_gather_chunk<<<16, 256, 0, c->stream>>>(A->d_volume, cx * 16, cy * 16, cz * 16,
	                                                  reinterpret_cast<float2 *>(A->d_chunkOut));
	c->launches++;
	OSB_CUDA(cudaMemcpyAsync(out, A->d_chunkOut, sizeof(float2) * 4096, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_ambient_evaluate_box(osb200_ctx *c, int x0, int y0, int z0, int x1, int y1, int z1, float *out) {
	if (!c || !c->mapLoaded || !out || x1 < x0 || y1 < y0 || z1 < z0) {
		osb::set_error("osb200_ambient_evaluate_box: bad argument or no map");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::ambient_init(c))
		return -2;
	const int nx = x1 - x0 + 1, ny = y1 - y0 + 1, nz = z1 - z0 + 1;
	const size_t nvox = (size_t)nx * ny * nz;
	if (c->ensure_scratch(nvox * 4))
		return -2;
	const size_t threads = nvox * 16;
	osb::k_ao_evaluate_box<<<(unsigned)((threads + 255) / 256), 256, 0, c->stream>>>(
	  c->d_solid, c->ambient->d_rays, x0, y0, z0, nx, ny, nz, reinterpret_cast<float *>(c->d_scratch));
	c->launches++;
	OSB_CUDA(cudaGetLastError());
	OSB_CUDA(cudaMemcpyAsync(out, c->d_scratch, nvox * 4, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_ambient_volume(osb200_ctx *c, const float **d_volume) {
	if (!c || !d_volume) {
		osb::set_error("osb200_ambient_volume: null argument");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::ambient_init(c))
		return -2;
	*d_volume = reinterpret_cast<const float *>(c->ambient->d_volume);
	return 0;
}

} // extern "C"
