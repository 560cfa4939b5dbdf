// Sun-shadow bitmap + radiosity on the GPU -- the CPU halves of the reference's OpenGL renderer:
//   GLMapShadowRenderer::GeneratePixel / BuildPixel        Sources/Draw/GLMapShadowRenderer.cpp:177-218
//   GLMapShadowRenderer::GameMapChanged / MarkUpdate        :220-229
//   GLMapShadowRenderer::Update (regenerate + invalidate)   :78-108
//   GLRadiosityRenderer::Evaluate                           Sources/Draw/GLRadiosityRenderer.cpp:162-275
//   GLRadiosityRenderer::GameMapChanged / Invalidate        :277-336
//   GLRadiosityRenderer::UpdateDirtyChunks (near-eye first) :394-450
//   GLRadiosityRenderer::EncodeValue / CompressDynamicRange :456-515
//   GLRadiosityRenderer::UpdateChunk                        :517-543
//
// Layout: the 512x512 shadow bitmap ([y][x], 1 MiB) and the four 512x512x64 RGB10A2 volumes
// ([z][y][x], 64 MiB each -- the layout of the reference's 3-D textures, GLRadiosityRenderer.cpp:120-129)
// stay resident; per-chunk dirty boxes live on the device as {min[3], max[3]} with "clean" = empty box,
// so the reference's "first invalidation assigns, later ones union" (:318-334) is plain atomicMin/atomicMax.
// One CTA evaluates one dirty 16^3 chunk: the (16+12) x (16+15+12) bitmap window every voxel of the chunk
// taps (centre = (x, y - z)) is staged in shared memory; one thread owns one voxel and walks its 13x13
// taps in the reference's order (x outer, y inner) with separate IEEE mul/add/div/sqrt (--fmad=false), so
// the 12 result floats and the 4 encoded words are bit-identical to the reference, not just within 1e-4.
#include "ctx.h"

#include <algorithm>
#include <cstring>

namespace osb {

	constexpr int kEnv = 6;       // GLRadiosityRenderer.h:44 Envelope
	constexpr int kChunk = 16;    // ChunkSize
	constexpr int kChunksXY = 32; // 512 / 16
	constexpr int kChunksZ = 4;
	constexpr int kNumChunks = kChunksXY * kChunksXY * kChunksZ;
	constexpr int kTileW = kChunk + 2 * kEnv;              // 28
	constexpr int kTileH = kChunk + (kChunk - 1) + 2 * kEnv; // 43
	constexpr uint32_t kInitWord = 0x20080200u;            // GLRadiosityRenderer.cpp:79-88

	// chunk index as the reference's GetChunk: (cx + cy * chunkW) * chunkD + cz (GLRadiosityRenderer.h:64-72)
	__host__ __device__ inline int chunk_index(int cx, int cy, int cz) { return (cx + cy * kChunksXY) * kChunksZ + cz; }

	struct ChunkBox { // clean <=> empty box (lo = 16, hi = -1)
		int lo[3], hi[3];
		int pad[2];
	};

	struct RadiosityState {
		uint32_t *d_bitmap = nullptr;     // [y][x]
		uint32_t *d_vol[4] = {nullptr, nullptr, nullptr, nullptr}; // flat, x, y, z: [z][y][x]
		ChunkBox *d_boxes = nullptr;      // kNumChunks
		int *d_list = nullptr;            // selected chunk ids + [kNumChunks] = count
		uint32_t *d_words = nullptr;      // dirty word list for the shadow update
		float2 *d_tapTab = nullptr;       // {1 / len, 1 / (len + 0.4)} of a tap by 4 * len^2 (k_rad_tap_table)
		std::vector<uint32_t> markWords;  // host dirty bitset, 16 words per row (GLMapShadowRenderer.cpp:57-58)
		bool allDirty = true;             // initial state: every pixel + every chunk dirty
		bool anyMarked = false;
		int *h_count = nullptr;           // pinned
		uint32_t *h_words = nullptr;      // pinned staging of the word list
	};

	// ---- shadow bitmap ---------------------------------------------------------------------------

	__device__ __forceinline__ uint32_t shadow_pack(int distance, uint32_t col, bool side) {
		const uint32_t r = (col & 0xff) >> 2, g = ((col >> 8) & 0xff) >> 2, b = ((col >> 16) & 0xff) >> 2;
		return r + (g << 8) + (b << 16) + ((uint32_t)distance << 24) + (side ? 0x80u : 0u);
	}

	// GeneratePixel: the diagonal walk visits column (x, y+k) for the y-plane test of z = k-1 and then the
	// z-plane test of z = k, so one 8-byte load per step serves both (65 loads instead of 128).
	__device__ uint32_t shadow_pixel(const uint64_t *__restrict__ solid, const uint32_t *__restrict__ color, int x,
	                                 int y) {
		for (int k = 0; k <= 64; k++) {
			const uint32_t t = tiled_index(x, (y + k) & 511);
			const uint64_t w = solid[t];
			if (k >= 1 && k - 1 < 63 && ((w >> (k - 1)) & 1ull)) // y-plane hit of z = k-1
				return shadow_pack(k, color[(size_t)t * 64 + (k - 1)], true);
			if (k < 63 && ((w >> k) & 1ull)) // z-plane hit of z = k
				return shadow_pack(k, color[(size_t)t * 64 + k], false);
		}
		return shadow_pack(64, color[(size_t)tiled_index(x, (y + 64) & 511) * 64 + 63], false);
	}

	// Invalidate(+-Envelope box around (x,y,z)), GLRadiosityRenderer.cpp:277-336
	__device__ void invalidate_around(ChunkBox *boxes, int x, int y, int z) {
		const int mn[3] = {x - kEnv, y - kEnv, max(z - kEnv, 0)};
		const int mx[3] = {x + kEnv, y + kEnv, min(z + kEnv, 63)};
		if (mn[2] > mx[2])
			return;
		for (int cx = mn[0] >> 4; cx <= (mx[0] >> 4); cx++)
			for (int cy = mn[1] >> 4; cy <= (mx[1] >> 4); cy++)
				for (int cz = mn[2] >> 4; cz <= (mx[2] >> 4); cz++) {
					ChunkBox &b = boxes[chunk_index(cx & 31, cy & 31, cz)];
					const int org[3] = {cx * 16, cy * 16, cz * 16};
#pragma unroll
					for (int k = 0; k < 3; k++) {
						atomicMin(&b.lo[k], max(mn[k] - org[k], 0));
						atomicMax(&b.hi[k], min(mx[k] - org[k], 15));
					}
				}
	}

	// GLMapShadowRenderer::Update: one warp per dirty 32-pixel word (words == nullptr: every word).
	__global__ void k_shadow_update(const uint64_t *__restrict__ solid, const uint32_t *__restrict__ color,
	                                uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ words, int nWords,
	                                ChunkBox *boxes, int invalidate) {
		const int wi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
		if (wi >= nWords)
			return;
		const uint32_t word = words ? words[wi] : (uint32_t)wi;
		const int y = word >> 4, x = ((word & 15) << 5) + (threadIdx.x & 31);
		const uint32_t nv = shadow_pixel(solid, color, x, y);
		uint32_t *p = bitmap + (size_t)y * 512 + x;
		const uint32_t ov = *p;
		if (ov != nv) {
			*p = nv;
			if (invalidate) {
				int dist = (int)(nv >> 24);
				invalidate_around(boxes, x, (y + dist) & 511, dist);
				dist = (int)(ov >> 24);
				invalidate_around(boxes, x, (y + dist) & 511, dist);
			}
		}
	}

	// ---- radiosity -------------------------------------------------------------------------------

	// CompressDynamicRange + one channel of EncodeValue (GLRadiosityRenderer.cpp:456-515)
	__device__ __forceinline__ uint32_t encode_channel(float c, int mode) {
		if (mode < 2)
			c = c >= 0.f ? __fsqrt_rn(c) : -__fsqrt_rn(-c);
		c = __fmul_rn(c, .5f);
		c = __fadd_rn(c, .5f);
		c = __fmul_rn(c, 1022.f / 1023.f);
		float v = __fadd_rn(__fmul_rn(c, 1023.f), .5f);
		if (v > 1023.2f)
			v = 1023.2f;
		if (v < 0.f)
			v = 0.f;
		int iv = (int)(unsigned int)__float2uint_rz(v);
		iv = min(max(iv, 0), 1023);
		return (uint32_t)iv;
	}
	__device__ __forceinline__ uint32_t encode_value(const float *v, int mode) {
		return 0xC0000000u | (encode_channel(v[0], mode) << 20) | (encode_channel(v[1], mode) << 10) |
		       encode_channel(v[2], mode);
	}

	// Evaluate for voxel (vx,vy,vz).  tile = bitmap window in shared memory, tile(0,0) = bitmap pixel
	// (tx0, ty0) (unwrapped coordinates; wrapped when the tile was loaded), row pitch kTileW.
	// The distance terms of a tap without the square root and the two divisions.  A tap's offset (fx, fy, fz) is made of
	// two integers and one half-integer (voxel centres against the face a bitmap pixel stands for), so 4 * len^2 is an
	// integer n, computed exactly, and {1 / len, 1 / (len + 0.4)} are functions of n alone: a table, filled once with the
	// very same IEEE operations (k_rad_tap_table), gives the same bits for 1 load instead of ~55 instructions.  n beyond
	// the table (a depth byte no valid bitmap holds) takes the arithmetic path.
	constexpr int kTapTabMax = 36863; // 4 * (6^2 + 69.5^2 + 64^2) = 35849 for depth bytes <= 64; 288 KB of float2
	__global__ void k_rad_tap_table(float2 *tab) {
		const int n = blockIdx.x * blockDim.x + threadIdx.x;
		if (n > kTapTabMax)
			return;
		const float len = __fsqrt_rn(__fmul_rn((float)n, 0.25f));
		tab[n] = make_float2(__fdiv_rn(1.f, len), __fdiv_rn(1.f, __fadd_rn(len, .4f)));
	}

	__device__ __forceinline__ void evaluate_voxel(const uint32_t *tile, int tx0, int ty0, int vx, int vy, int vz,
	                                               float acc[12], const float2 *__restrict__ tapTab) {
#pragma unroll
		for (int i = 0; i < 12; i++)
			acc[i] = 0.f;
		const float px = __fadd_rn((float)vx, .5f), py = __fadd_rn((float)vy, .5f), pz = __fadd_rn((float)vz, .5f);
		const int cx = vx, cy = vy - vz;
		const uint32_t *base = tile + (cy - kEnv - ty0) * kTileW + (cx - kEnv - tx0);
#pragma unroll 1
		for (int dx = -kEnv; dx <= kEnv; dx++) {
			const float fx = __fsub_rn(px, __fadd_rn((float)(cx + dx), .5f));
			const float fx2 = __fmul_rn(fx, fx);
#pragma unroll 1
			for (int dy = -kEnv; dy <= kEnv; dy++) {
				const uint32_t pixel = base[(dy + kEnv) * kTileW + (dx + kEnv)];
				const int depth = (int)(pixel >> 24);
				const int wy = cy + dy + depth, wz = depth;
				float fy, fz, facing;
				int a2, b2; // 2 fy, 2 fz
				if (pixel & 0x80u) {
					if (wy <= vy)
						continue;
					fy = __fsub_rn(py, (float)wy);
					fz = __fsub_rn(pz, __fsub_rn((float)wz, .5f));
					facing = -fy;
					a2 = 2 * (vy - wy) + 1;
					b2 = 2 * (vz - wz + 1);
				} else {
					if (wz <= vz)
						continue;
					fy = __fsub_rn(py, __fadd_rn((float)wy, .5f));
					fz = __fsub_rn(pz, (float)wz);
					facing = -fz;
					a2 = 2 * (vy - wy);
					b2 = 2 * (vz - wz) + 1;
				}
				// 4 * len^2 in integers: (2 fx)^2 + (2 fy)^2 + (2 fz)^2 (every term of the float sum is exact as well)
				const int n = 4 * dx * dx + a2 * a2 + b2 * b2;
				float invLen, invSmooth;
				if (n <= kTapTabMax) {
					const float2 t = __ldg(tapTab + n);
					invLen = t.x;
					invSmooth = t.y;
				} else {
					const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(fx2, __fmul_rn(fy, fy)), __fmul_rn(fz, fz)));
					invLen = __fdiv_rn(1.f, len);
					invSmooth = __fdiv_rn(1.f, __fadd_rn(len, .4f));
				}
				float intensity = __fmul_rn(facing, invLen);
				intensity = __fmul_rn(intensity, invSmooth);
				intensity = __fmul_rn(intensity, invSmooth);
				const float s = -invLen;
				const float nx = __fmul_rn(fx, s), ny = __fmul_rn(fy, s), nz = __fmul_rn(fz, s);
				float c[3];
				c[0] = __fmul_rn((float)(pixel & 0x3f), intensity);
				c[1] = __fmul_rn((float)((pixel >> 8) & 0x3f), intensity);
				c[2] = __fmul_rn((float)((pixel >> 16) & 0x3f), intensity);
#pragma unroll
				for (int k = 0; k < 3; k++) {
					acc[k] = __fadd_rn(acc[k], c[k]);
					acc[3 + k] = __fadd_rn(acc[3 + k], __fmul_rn(c[k], nx));
					acc[6 + k] = __fadd_rn(acc[6 + k], __fmul_rn(c[k], ny));
					acc[9 + k] = __fadd_rn(acc[9 + k], __fmul_rn(c[k], nz));
				}
			}
		}
		const float scale = 0.1f / 64.f;
#pragma unroll
		for (int i = 0; i < 12; i++)
			acc[i] = __fmul_rn(acc[i], scale);
	}

	// loads the bitmap window of the box [x0..x1] x [y0..y1] x [z0..z1] (spans <= 16 per axis)
	__device__ __forceinline__ void load_tile(uint32_t *tile, const uint32_t *__restrict__ bitmap, int tx0, int ty0) {
		for (int i = threadIdx.x; i < kTileW * kTileH; i += blockDim.x) {
			const int r = i / kTileW, c = i - r * kTileW;
			tile[i] = bitmap[(size_t)((ty0 + r) & 511) * 512 + ((tx0 + c) & 511)];
		}
	}

	// UpdateDirtyChunks' candidate selection (:394-430) without its random pick: dirty chunks within 6 chunks
	// of the eye first; if there are none, any dirty chunk; at most maxChunks of them, in index order.
	// One CTA of 1024 threads; list[kNumChunks] receives the count.
	__global__ void __launch_bounds__(1024) k_rad_select(const ChunkBox *__restrict__ boxes, int *list, int maxChunks,
	                                                     int useEye, int eyeX, int eyeY, int eyeZ) {
		__shared__ int s_cnt[2];
		__shared__ int s_warp[32];
		if (threadIdx.x < 2)
			s_cnt[threadIdx.x] = 0;
		__syncthreads();
		for (int pass = 0; pass < 2; pass++) {
			// pass 0: near chunks (only with an eye); pass 1: all chunks, only if pass 0 found nothing
			if (pass == 0 && !useEye)
				continue;
			if (pass == 1 && s_cnt[0] > 0)
				break;
			for (int base = 0; base < kNumChunks; base += 1024) {
				const int i = base + threadIdx.x;
				const ChunkBox b = boxes[i];
				bool take = b.hi[0] >= b.lo[0] && b.hi[1] >= b.lo[1] && b.hi[2] >= b.lo[2];
				if (take && pass == 0) {
					const int cz = i & 3, cx = (i >> 2) & 31, cy = i >> 7;
					const int dx = (cx - eyeX) & 31, dy = (cy - eyeY) & 31, dz = cz - eyeZ;
					if ((dx >= 6 && dx <= 32 - 6) || (dy >= 6 && dy <= 32 - 6) || dz >= 6 || dz <= -6)
						take = false;
				}
				// block-wide ordered compaction
				const unsigned bal = __ballot_sync(0xffffffffu, take);
				const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
				if (lane == 0)
					s_warp[warp] = __popc(bal);
				__syncthreads();
				int off = s_cnt[1];
				for (int w = 0; w < warp; w++)
					off += s_warp[w];
				off += __popc(bal & ((1u << lane) - 1u));
				if (take && off < maxChunks)
					list[off] = i;
				__syncthreads();
				if (threadIdx.x == 0) {
					int tot = 0;
					for (int w = 0; w < 32; w++)
						tot += s_warp[w];
					s_cnt[1] += tot;
				}
				__syncthreads();
			}
			if (threadIdx.x == 0)
				s_cnt[0] = s_cnt[1];
			__syncthreads();
		}
		if (threadIdx.x == 0)
			list[kNumChunks] = min(s_cnt[1], maxChunks);
	}

	// UpdateChunk: one CTA per selected chunk, one thread per voxel of its dirty box (strided).
	__global__ void __launch_bounds__(256) k_radiosity(const uint32_t *__restrict__ bitmap, ChunkBox *boxes,
	                                                   const int *__restrict__ list, uint32_t *__restrict__ vFlat,
	                                                   uint32_t *__restrict__ vX, uint32_t *__restrict__ vY,
	                                                   uint32_t *__restrict__ vZ, int mode, const float2 *__restrict__ tapTab) {
		__shared__ uint32_t tile[kTileW * kTileH];
		__shared__ ChunkBox s_box;
		if ((int)blockIdx.x >= list[kNumChunks])
			return;
		const int ci = list[blockIdx.x];
		if (threadIdx.x == 0)
			s_box = boxes[ci];
		const int cz = ci & 3, cx = (ci >> 2) & 31, cy = ci >> 7;
		const int X0 = cx * 16, Y0 = cy * 16, Z0 = cz * 16;
		const int tx0 = X0 - kEnv, ty0 = Y0 - Z0 - (kChunk - 1) - kEnv;
		load_tile(tile, bitmap, tx0, ty0);
		__syncthreads();
		const ChunkBox b = s_box;
		const int nx = b.hi[0] - b.lo[0] + 1, ny = b.hi[1] - b.lo[1] + 1, nz = b.hi[2] - b.lo[2] + 1;
		const int n = nx * ny * nz;
		for (int i = threadIdx.x; i < n; i += blockDim.x) {
			const int lx = i % nx, t = i / nx, ly = t % ny, lz = t / ny;
			const int x = X0 + b.lo[0] + lx, y = Y0 + b.lo[1] + ly, z = Z0 + b.lo[2] + lz;
			float acc[12];
			evaluate_voxel(tile, tx0, ty0, x, y, z, acc, tapTab);
			const size_t o = ((size_t)z * 512 + y) * 512 + x;
			vFlat[o] = encode_value(acc, mode);
			vX[o] = encode_value(acc + 3, mode);
			vY[o] = encode_value(acc + 6, mode);
			vZ[o] = encode_value(acc + 9, mode);
		}
		if (threadIdx.x == 0) { // c.dirty = false
			ChunkBox clean;
			clean.lo[0] = clean.lo[1] = clean.lo[2] = 16;
			clean.hi[0] = clean.hi[1] = clean.hi[2] = -1;
			clean.pad[0] = clean.pad[1] = 0;
			boxes[ci] = clean;
		}
	}

	// Test hook: un-encoded Evaluate results of an arbitrary box (<= 16 per axis per CTA), 12 floats/voxel,
	// box-ordered [z][y][x].  Grid = blocks of 16^3 covering the box.
	__global__ void __launch_bounds__(256) k_radiosity_f32(const uint32_t *__restrict__ bitmap, int x0, int y0, int z0,
	                                                       int nx, int ny, int nz, float *__restrict__ out,
	                                                       const float2 *__restrict__ tapTab) {
		__shared__ uint32_t tile[kTileW * kTileH];
		const int bxn = (nx + 15) >> 4, byn = (ny + 15) >> 4;
		const int bx = blockIdx.x % bxn, by = (blockIdx.x / bxn) % byn, bz = blockIdx.x / (bxn * byn);
		const int X0 = x0 + bx * 16, Y0 = y0 + by * 16, Z0 = z0 + bz * 16;
		const int tx0 = X0 - kEnv, ty0 = Y0 - Z0 - (kChunk - 1) - kEnv;
		load_tile(tile, bitmap, tx0, ty0);
		__syncthreads();
		for (int i = threadIdx.x; i < 4096; i += blockDim.x) {
			const int lx = i & 15, ly = (i >> 4) & 15, lz = i >> 8;
			const int gx = bx * 16 + lx, gy = by * 16 + ly, gz = bz * 16 + lz;
			if (gx >= nx || gy >= ny || gz >= nz)
				continue;
			float acc[12];
			evaluate_voxel(tile, tx0, ty0, X0 + lx, Y0 + ly, Z0 + lz, acc, tapTab);
			float *o = out + (((size_t)gz * ny + gy) * nx + gx) * 12;
#pragma unroll
			for (int k = 0; k < 12; k++)
				o[k] = acc[k];
		}
	}

	// Measurement hook: how many of the 169 taps per voxel pass Evaluate's normal cull (GLRadiosityRenderer.cpp:199-214) and
	// do the float work -- the executed-tap count behind bench.py's FP32 figure for C5.  One CTA per 16^3 chunk.
	__global__ void __launch_bounds__(256) k_radiosity_count_taps(const uint32_t *__restrict__ bitmap, unsigned long long *out) {
		__shared__ uint32_t tile[kTileW * kTileH];
		__shared__ unsigned int s_sum;
		const int ci = blockIdx.x;
		const int cz = ci & 3, cx = (ci >> 2) & 31, cy = ci >> 7;
		const int X0 = cx * 16, Y0 = cy * 16, Z0 = cz * 16;
		const int tx0 = X0 - kEnv, ty0 = Y0 - Z0 - (kChunk - 1) - kEnv;
		if (threadIdx.x == 0)
			s_sum = 0;
		load_tile(tile, bitmap, tx0, ty0);
		__syncthreads();
		unsigned int mine = 0;
		for (int i = threadIdx.x; i < 4096; i += blockDim.x) {
			const int vx = X0 + (i & 15), vy = Y0 + ((i >> 4) & 15), vz = Z0 + (i >> 8);
			const int ccx = vx, ccy = vy - vz;
			const uint32_t *base = tile + (ccy - kEnv - ty0) * kTileW + (ccx - kEnv - tx0);
			for (int dx = -kEnv; dx <= kEnv; dx++)
				for (int dy = -kEnv; dy <= kEnv; dy++) {
					const uint32_t pixel = base[(dy + kEnv) * kTileW + (dx + kEnv)];
					const int depth = (int)(pixel >> 24);
					const int wy = ccy + dy + depth, wz = depth;
					mine += (pixel & 0x80u) ? (wy > vy) : (wz > vz);
				}
		}
		atomicAdd(&s_sum, mine);
		__syncthreads();
		if (threadIdx.x == 0)
			atomicAdd(out, (unsigned long long)s_sum);
	}

	__global__ void k_fill_u32(uint32_t *p, size_t n, uint32_t v) {
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
			p[i] = v;
	}
	__global__ void k_reset_boxes(ChunkBox *b, int n) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i < n) {
			b[i].lo[0] = b[i].lo[1] = b[i].lo[2] = 0;
			b[i].hi[0] = b[i].hi[1] = b[i].hi[2] = 15;
			b[i].pad[0] = b[i].pad[1] = 0;
		}
	}
	__global__ void k_gather_chunk(const uint32_t *__restrict__ v0, const uint32_t *__restrict__ v1,
	                               const uint32_t *__restrict__ v2, const uint32_t *__restrict__ v3, int X0, int Y0,
	                               int Z0, uint32_t *__restrict__ out) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x; // 0..4095, [z][y][x]
		const int lx = i & 15, ly = (i >> 4) & 15, lz = i >> 8;
		const size_t o = ((size_t)(Z0 + lz) * 512 + (Y0 + ly)) * 512 + (X0 + lx);
		out[i] = v0[o];
		out[4096 + i] = v1[o];
		out[8192 + i] = v2[o];
		out[12288 + i] = v3[o];
	}

	// ---- host side -------------------------------------------------------------------------------

	static int radiosity_reset(osb200_ctx *c) {
		RadiosityState *R = c->radiosity;
		// state of freshly constructed GLMapShadowRenderer / GLRadiosityRenderer: bitmap 0xffffffff with every
		// word dirty (GLMapShadowRenderer.cpp:62-66), volumes 0x20080200, every chunk dirty with a full box
		OSB_CUDA(cudaMemsetAsync(R->d_bitmap, 0xff, 512 * 512 * 4, c->stream));
		for (int i = 0; i < 4; i++)
			k_fill_u32<<<c->smCount * 8, 256, 0, c->stream>>>(R->d_vol[i], (size_t)512 * 512 * 64, kInitWord);
		k_reset_boxes<<<kNumChunks / 256, 256, 0, c->stream>>>(R->d_boxes, kNumChunks);
		c->launches += 5;
		std::fill(R->markWords.begin(), R->markWords.end(), 0u);
		R->allDirty = true;
		R->anyMarked = false;
		OSB_CUDA(cudaGetLastError());
		return 0;
	}

	static int radiosity_init(osb200_ctx *c) {
		if (c->radiosity)
			return 0;
		auto *R = new RadiosityState();
		c->radiosity = R;
		R->markWords.assign(512 * 16, 0u);
		OSB_CUDA(cudaMalloc(&R->d_bitmap, 512 * 512 * 4));
		for (int i = 0; i < 4; i++)
			OSB_CUDA(cudaMalloc(&R->d_vol[i], (size_t)512 * 512 * 64 * 4));
		OSB_CUDA(cudaMalloc(&R->d_boxes, sizeof(ChunkBox) * kNumChunks));
		OSB_CUDA(cudaMalloc(&R->d_list, sizeof(int) * (kNumChunks + 1)));
		OSB_CUDA(cudaMalloc(&R->d_words, 512 * 16 * 4));
		OSB_CUDA(cudaMalloc(&R->d_tapTab, sizeof(float2) * (kTapTabMax + 1)));
		k_rad_tap_table<<<(kTapTabMax + 256) / 256, 256, 0, c->stream>>>(R->d_tapTab);
		c->launches++;
		OSB_CUDA(cudaMallocHost(&R->h_count, sizeof(int)));
		OSB_CUDA(cudaMallocHost(&R->h_words, 512 * 16 * 4));
		return radiosity_reset(c);
	}

	void radiosity_destroy(osb200_ctx *c) {
		RadiosityState *R = c->radiosity;
		if (!R)
			return;
		cudaFree(R->d_bitmap);
		for (int i = 0; i < 4; i++)
			cudaFree(R->d_vol[i]);
		cudaFree(R->d_boxes);
		cudaFree(R->d_list);
		cudaFree(R->d_words);
		cudaFree(R->d_tapTab);
		cudaFreeHost(R->h_count);
		cudaFreeHost(R->h_words);
		delete R;
		c->radiosity = nullptr;
	}

	// osb200_map_upload replaced the map: same as constructing new shadow/radiosity renderers for it
	void radiosity_map_replaced(osb200_ctx *c) {
		if (c->radiosity)
			radiosity_reset(c);
	}

	// GLMapShadowRenderer::GameMapChanged -> MarkUpdate(x, y - z), MarkUpdate(x, y - z - 1) (:220-229)
	void radiosity_mark_dirty(osb200_ctx *c, int x, int y, int z) {
		RadiosityState *R = c->radiosity;
		if (!R || R->allDirty)
			return; // not in use yet, or everything is dirty anyway
		for (int k = 0; k < 2; k++) {
			const int px = x & 511, py = (y - z - k) & 511;
			R->markWords[(px >> 5) + py * 16] |= 1u << (px & 31);
		}
		R->anyMarked = true;
	}

	// GLMapShadowRenderer::Update for the pending marks.  The device map must already hold the edits.
	static int shadow_sync(osb200_ctx *c) {
		if (radiosity_init(c))
			return -2;
		RadiosityState *R = c->radiosity;
		if (c->anyDirty) {
			set_error("shadow/radiosity update: edited columns are pending -- call osb200_map_flush first");
			return -1;
		}
		if (R->allDirty) {
			// every chunk box is already full, so the per-pixel invalidation cannot change anything: skipped
			KLaunch kl(c, K_SHADOW_BITMAP);
			k_shadow_update<<<512 * 16 / 8, 256, 0, c->stream>>>(c->d_solid, c->d_color, R->d_bitmap, nullptr,
			                                                    512 * 16, R->d_boxes, 0);
			R->allDirty = false;
		} else if (R->anyMarked) {
			OSB_CUDA(cudaStreamSynchronize(c->stream)); // h_words may still be in flight
			int n = 0;
			for (int i = 0; i < 512 * 16; i++)
				if (R->markWords[i]) {
					R->h_words[n++] = (uint32_t)i;
					R->markWords[i] = 0;
				}
			R->anyMarked = false;
			OSB_CUDA(cudaMemcpyAsync(R->d_words, R->h_words, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
			KLaunch kl(c, K_SHADOW_BITMAP);
			k_shadow_update<<<(n + 7) / 8, 256, 0, c->stream>>>(c->d_solid, c->d_color, R->d_bitmap, R->d_words, n,
			                                                   R->d_boxes, 1);
		}
		OSB_CUDA(cudaGetLastError());
		return 0;
	}
} // namespace osb

extern "C" {

int osb200_shadow_bitmap(osb200_ctx *c, uint32_t *out) {
	if (!c || !c->mapLoaded) {
		osb::set_error("osb200_shadow_bitmap: no context or no map uploaded");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	int rc = osb::shadow_sync(c);
	if (rc)
		return rc;
	if (out)
		OSB_CUDA(cudaMemcpyAsync(out, c->radiosity->d_bitmap, 512 * 512 * 4, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_radiosity_update(osb200_ctx *c, int maxChunks, const float *eye) {
	if (!c || !c->mapLoaded) {
		osb::set_error("osb200_radiosity_update: no context or no map uploaded");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	int rc = osb::shadow_sync(c);
	if (rc)
		return rc;
	osb::RadiosityState *R = c->radiosity;
	if (maxChunks < 0 || maxChunks >= osb::kNumChunks) {
		maxChunks = osb::kNumChunks;
		eye = nullptr; // "all dirty chunks": the near-eye priority only orders a budgeted update
	}
	if (maxChunks == 0)
		return 0;
	int ex = 0, ey = 0, ez = 0;
	if (eye) { // (int)eyePos >> ChunkSizeBits, GLRadiosityRenderer.cpp:401-404
		ex = (int)eye[0] >> 4;
		ey = (int)eye[1] >> 4;
		ez = (int)eye[2] >> 4;
	}
	{
		osb::KLaunch kl(c, osb::K_RAD_SELECT);
		osb::k_rad_select<<<1, 1024, 0, c->stream>>>(R->d_boxes, R->d_list, maxChunks, eye ? 1 : 0, ex, ey, ez);
	}
	{
		osb::KLaunch kl(c, osb::K_RADIOSITY);
		osb::k_radiosity<<<maxChunks, 256, 0, c->stream>>>(R->d_bitmap, R->d_boxes, R->d_list, R->d_vol[0], R->d_vol[1],
		                                                  R->d_vol[2], R->d_vol[3], c->opts.radiosity_mode, R->d_tapTab);
	}
	OSB_CUDA(cudaGetLastError());
	OSB_CUDA(cudaMemcpyAsync(R->h_count, R->d_list + osb::kNumChunks, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return *R->h_count;
}

int osb200_radiosity_chunk_info(osb200_ctx *c, int cx, int cy, int cz, int out[7]) {
	if (!c || !c->mapLoaded || !out || cx < 0 || cx >= 32 || cy < 0 || cy >= 32 || cz < 0 || cz >= 4) {
		osb::set_error("osb200_radiosity_chunk_info: bad argument or no map");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	int rc = osb::shadow_sync(c);
	if (rc)
		return rc;
	osb::ChunkBox b;
	OSB_CUDA(cudaMemcpyAsync(&b, c->radiosity->d_boxes + osb::chunk_index(cx, cy, cz), sizeof(b),
	                         cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	out[0] = (b.hi[0] >= b.lo[0] && b.hi[1] >= b.lo[1] && b.hi[2] >= b.lo[2]) ? 1 : 0;
	for (int k = 0; k < 3; k++) {
		out[1 + k] = b.lo[k];
		out[4 + k] = b.hi[k];
	}
	return 0;
}

int osb200_radiosity_read_chunk(osb200_ctx *c, int cx, int cy, int cz, uint32_t *flat, uint32_t *x, uint32_t *y,
                                uint32_t *z) {
	if (!c || !c->mapLoaded || cx < 0 || cx >= 32 || cy < 0 || cy >= 32 || cz < 0 || cz >= 4) {
		osb::set_error("osb200_radiosity_read_chunk: bad argument or no map");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::radiosity_init(c))
		return -2;
	osb::RadiosityState *R = c->radiosity;
	if (c->ensure_scratch(4 * 4096 * 4))
		return -2;
	uint32_t *tmp = reinterpret_cast<uint32_t *>(c->d_scratch);
	osb::k_gather_chunk<<<16, 256, 0, c->stream>>>(R->d_vol[0], R->d_vol[1], R->d_vol[2], R->d_vol[3], cx * 16, cy * 16,
	                                              cz * 16, tmp);
	c->launches++;
	uint32_t *dst[4] = {flat, x, y, z};
	for (int i = 0; i < 4; i++)
		if (dst[i])
			OSB_CUDA(cudaMemcpyAsync(dst[i], tmp + i * 4096, 4096 * 4, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_radiosity_evaluate_box(osb200_ctx *c, int x0, int y0, int z0, int x1, int y1, int z1, float *out) {
	if (!c || !c->mapLoaded || !out || x1 < x0 || y1 < y0 || z1 < z0) {
		osb::set_error("osb200_radiosity_evaluate_box: bad argument or no map");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	int rc = osb::shadow_sync(c);
	if (rc)
		return rc;
	const int nx = x1 - x0 + 1, ny = y1 - y0 + 1, nz = z1 - z0 + 1;
	const size_t bytes = (size_t)nx * ny * nz * 12 * 4;
	if (c->ensure_scratch(bytes))
		return -2;
	const int blocks = ((nx + 15) >> 4) * ((ny + 15) >> 4) * ((nz + 15) >> 4);
	osb::k_radiosity_f32<<<blocks, 256, 0, c->stream>>>(c->radiosity->d_bitmap, x0, y0, z0, nx, ny, nz,
	                                                   reinterpret_cast<float *>(c->d_scratch), c->radiosity->d_tapTab);
	c->launches++;
	OSB_CUDA(cudaGetLastError());
	OSB_CUDA(cudaMemcpyAsync(out, c->d_scratch, bytes, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_radiosity_count_taps(osb200_ctx *c, unsigned long long *executed, unsigned long long *total) {
	if (!c || !c->mapLoaded || !executed) {
		osb::set_error("osb200_radiosity_count_taps: bad argument or no map");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	int rc = osb::shadow_sync(c);
	if (rc)
		return rc;
	if (c->ensure_scratch(sizeof(unsigned long long)))
		return -2;
	OSB_CUDA(cudaMemsetAsync(c->d_scratch, 0, sizeof(unsigned long long), c->stream));
	osb::k_radiosity_count_taps<<<osb::kNumChunks, 256, 0, c->stream>>>(c->radiosity->d_bitmap,
	                                                                  reinterpret_cast<unsigned long long *>(c->d_scratch));
	c->launches++;
	OSB_CUDA(cudaGetLastError());
	OSB_CUDA(cudaMemcpyAsync(executed, c->d_scratch, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	if (total)
		*total = 512ull * 512ull * 64ull * 169ull;
	return 0;
}

int osb200_radiosity_volumes(osb200_ctx *c, const uint32_t **flat, const uint32_t **x, const uint32_t **y,
                             const uint32_t **z, const uint32_t **shadow) {
	if (!c) {
		osb::set_error("osb200_radiosity_volumes: null context");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::radiosity_init(c))
		return -2;
	osb::RadiosityState *R = c->radiosity;
	if (flat)
		*flat = R->d_vol[0];
	if (x)
		*x = R->d_vol[1];
	if (y)
		*y = R->d_vol[2];
	if (z)
		*z = R->d_vol[3];
	if (shadow)
		*shadow = R->d_bitmap;
	return 0;
}

} // extern "C"
