// Software-renderer terrain pass + fog on the GPU, reproducing the reference's results:
//   SWRenderer::EndScene / BuildFrustrum   Sources/Draw/SWRenderer.cpp:207-227,709-745
//   SWMapRenderer::RenderInner             Sources/Draw/SWMapRenderer.cpp:1045-1174  (per-view set-up)
//   SWMapRenderer::BuildLine               Sources/Draw/SWMapRenderer.cpp:219-671    (k_build_lines)
//   SWMapRenderer::RenderFinal             Sources/Draw/SWMapRenderer.cpp:744-1043   (k_gather_fog)
//   SWRenderer::ApplyFog                   Sources/Draw/SWRenderer.cpp:341-516       (fused into k_gather_fog)
//
// Design: the reference renders a view as ~2.5-5k vertical "yaw lines", each a 1-D strip
// of (colour, depth) built by a 2-D DDA through map columns, then gathers every screen pixel from
// (line, pitch index).  Here
//   * one thread builds one line; the six per-column RLE z-lists of the reference are never stored --
//     tops/bottoms/walls are bit expressions of the 64-bit column mask and of the mask of the cell the
//     DDA just left, so a step costs one 8-byte load from the 2 MiB tiled bitmap;
//   * "first writer wins" along a strip is kept in a per-thread bitmask in shared memory, so strip
//     pixels are written once, write-only, and colours are fetched only for spans that land;
//   * gather and fog are one kernel: each warp owns an 8-pixel-wide column of 8x8 blocks, replays the
//     reference's per-thread float recurrences for its start vectors and writes BGRX + depth once.
// All float arithmetic is separate IEEE mul/add (--fmad=false) in the reference's evaluation order.
#include "ctx.h"
#include "libm_f32.cuh"
#include "numerics.cuh"

#include <cmath>
#include <cstdlib>
#include <cstring>

namespace osb {

	// Per-view constants.  Everything that the reference computes once per frame with libm
	// (tanf/atanf/asinf/atan2l/cosf/sinf, SWMapRenderer.cpp:1057-1133, SWRenderer.cpp:207-227,345-357)
	// is evaluated on the host with the same libm and shipped here.
	struct ViewParams {
		float origin[3];
		float right[3], up[3], front[3];
		float frN[4][3]; // frustum side-plane normals, SWRenderer::frustrum[2..5].n
		float fovXt, fovYt;
		float yawMin, yawMax;
		float cosYawMin, sinYawMin, cosScl, sinScl;
		int numLines, lineRes;
		int yawScale2, yawMin2;
		uint32_t lineBase;  // first line slot of this view inside the chunk
		uint32_t tabBase;   // first float of this view's recurrence tables
		float fogScale;     // 255 / fogDistance
		float dvx, dvy;     // ApplyFog steps
		uint32_t pad;
	};

	// Recurrence tables of one view, laid out at tabs + tabBase:
	//   vx[fw/4]            ApplyFog per-4-pixel-column vx          (SWRenderer.cpp:372,409)
	//   vy[fh/4]            ApplyFog per-4-row vy, per emulated thread band (:365-369,412)
	//   v1[fw/8][3]         RenderFinal per-8-pixel-column start vector, per emulated thread (:757,789,1040)
	__host__ __device__ inline uint32_t tab_floats(int fw, int fh) { return fw / 4 + fh / 4 + (fw / 8) * 3; }

	struct FrameGeom {
		int fw, fh, under, swThreads, sse2Fog, sse2Shade;
		int noFog; // fog deferred to k_scene_finish (models / dynamic lights are drawn between terrain and fog, SWRenderer.cpp:727-744)
		int noMap; // no map set: SWRenderer::EndScene still clears and fogs over the stale depth (SWRenderer.cpp:713-744)
		uint32_t fogB, fogG, fogR;
	};

	// One sequential float recurrence per role; the same code runs on the host (few views) or on one
	// lane per role on the device (many views).
	__host__ __device__ inline void prep_line_dirs(const ViewParams &v, float2 *dirs) {
		// SWMapRenderer.cpp:1129-1143
		float hx = v.cosYawMin, hy = v.sinYawMin;
		const float c = v.cosScl, s = v.sinScl;
		for (int i = 0; i < v.numLines; i++) {
			dirs[i] = make_float2(hx, hy);
			const float x = osbm::subf(osbm::mulf(hx, c), osbm::mulf(hy, s));
			const float y = osbm::addf(osbm::mulf(hx, s), osbm::mulf(hy, c));
			hx = x;
			hy = y;
		}
	}
	__host__ __device__ inline void prep_fog_vx(const ViewParams &v, const FrameGeom &g, float *vx) {
		float x = v.fovXt;
		for (int i = 0; i < g.fw / 4; i++) {
			vx[i] = x;
			x = osbm::addf(x, v.dvx);
		}
	}
	__host__ __device__ inline void prep_fog_vy(const ViewParams &v, const FrameGeom &g, int t, float *vy) {
		// band of emulated thread t, SWRenderer.cpp:359-369
		int startY = (int)((unsigned)g.fh * (unsigned)t / (unsigned)g.swThreads) & ~3;
		int endY = (int)((unsigned)g.fh * (unsigned)(t + 1) / (unsigned)g.swThreads) & ~3;
		float y = osbm::addf(v.fovYt, osbm::mulf(v.dvy, (float)(startY >> 2)));
		for (int r = startY; r < endY; r += 4) {
			vy[r >> 2] = y;
			y = osbm::addf(y, v.dvy);
		}
	}
	__host__ __device__ inline void prep_v1(const ViewParams &v, const FrameGeom &g, int t, float *v1tab) {
		// SWMapRenderer.cpp:757-759,774-781,789,1040
		const float fx = v.fovXt, fy = v.fovYt;
		float v1[3], dR[3], dRL[3];
		const float sR = osbm::mulf(osbm::divf(osbm::mulf(fx, 2.f), (float)g.fw), (float)g.under);
		unsigned startX = (unsigned)t * (unsigned)g.fw / (unsigned)g.swThreads;
		unsigned endX = (unsigned)(t + 1) * (unsigned)g.fw / (unsigned)g.swThreads;
		startX = (startX / 8) * 8;
		endX = (endX / 8) * 8;
		for (int k = 0; k < 3; k++) {
			v1[k] = osbm::addf(osbm::subf(v.front[k], osbm::mulf(v.right[k], fx)), osbm::mulf(v.up[k], fy));
			dR[k] = osbm::mulf(v.right[k], sR);
			dRL[k] = osbm::mulf(dR[k], (float)(8 / g.under));
			v1[k] = osbm::addf(v1[k], osbm::mulf(dR[k], (float)(startX / (unsigned)g.under)));
		}
		for (unsigned x = startX; x < endX; x += 8) {
			for (int k = 0; k < 3; k++) {
				v1tab[(x >> 3) * 3 + k] = v1[k];
				v1[k] = osbm::addf(v1[k], dRL[k]);
			}
		}
	}

	// Per-chunk working set.  Two of them, each with its own stream: chunk k+1's line build runs beside chunk k's
	// gather and fills the SMs that chunk k's launch tail leaves idle (a single yaw line is ~150 dependent DDA steps,
	// so every k_build_lines launch ends in a long low-occupancy tail).
	// lower bound of lanes per warp for small k_build_lines launches (OSB200_LINE_MIN_LANES, default 4; 32 = off).
	// One 1080p frame (3.8 k lines): 32 / 16 / 8 / 4 / 2 lanes -> 0.338 / 0.285 / 0.240 / 0.207 / 0.203 ms
	static int line_min_lanes() {
		static const int v = [] {
			const char *e = getenv("OSB200_LINE_MIN_LANES");
			const int n = e ? atoi(e) : 4;
			return n >= 32 ? 32 : (n >= 16 ? 16 : (n >= 8 ? 8 : (n >= 4 ? 4 : (n >= 2 ? 2 : 1))));
		}();
		return v;
	}

	struct ChunkBufs {
		float2 *d_lineDirs = nullptr;
		int4 *d_lineParams = nullptr;
		uint2 *d_strips = nullptr;
		uint32_t *d_lineMask = nullptr; // lineRes bits per line: strip pixels the line wrote
		float *d_tabs = nullptr;
		size_t lineCap = 0, lineParamCap = 0, stripCap = 0, maskCap = 0, tabCap = 0;
		cudaStream_t stream = nullptr;
		cudaEvent_t done = nullptr; // last kernel of the newest chunk that used this set
	};

	struct TerrainState {
		ViewParams *d_views = nullptr;
		ViewParams *h_views = nullptr; // pinned
		size_t viewCap = 0, dViewCap = 0, hTabCap = 0;
		ChunkBufs set[2];
		float *h_tabs = nullptr; // pinned (host-prep path)
		float2 *h_lineDirs = nullptr;
		size_t hLineCap = 0;
		uint16_t *d_atan = nullptr; // 4 x 5000: sm, lg, smN, lgN
		// persistent output buffers of the host-pointer entry point
		uint32_t *d_fb = nullptr;
		float *d_depth = nullptr;
		size_t outCap = 0;
		cudaEvent_t stageFree = nullptr;
		cudaStream_t copyStream = nullptr; // D2H of finished chunks, overlapped with the next chunk's kernels
		cudaEvent_t fork = nullptr, copyDone = nullptr;
	};
} // namespace osb

namespace osb {

	#ifndef OSB_LINE_THREADS
#define OSB_LINE_THREADS 64
#endif
	constexpr int kLineThreads = OSB_LINE_THREADS; // lines per CTA of k_build_lines
#ifndef OSB_COLOR_EARLY
#define OSB_COLOR_EARLY 0
#endif
#ifndef OSB_EXP
#define OSB_EXP 0
#endif
#ifndef OSB_RCP_SMEM
#define OSB_RCP_SMEM 0 // measured on B200: no gain over the L1-cached global table (3.79 vs 3.76 ms), left as a knob
#endif
#ifndef OSB_TRANSFORM_BRANCHY
#define OSB_TRANSFORM_BRANCHY 1
#endif
#ifdef OSB_STATS
	__device__ unsigned long long g_stats[16];
#define STAT(i, n) atomicAdd(&g_stats[i], (unsigned long long)(n))
#else
#define STAT(i, n)
#endif
	enum { F_POSX = 0, F_NEGX = 1, F_POSY = 2, F_NEGY = 3, F_POSZ = 4, F_NEGZ = 5 }; // SWMapRenderer.cpp:68

	// BuildLinePixel colour (SWMapRenderer.cpp:532-576). col = 0xHHBBGGRR from the map; mode 0 = NegY walls and
	// tops (NegZ), 1 = PosX/PosY/NegX walls, 2 = undersides (PosZ).  All four 8-bit lanes are shaded in one
	// register: (x>>1)+(x>>2) <= 190 never carries into the neighbouring lane.
	template <bool kSse2Shade> __device__ __forceinline__ uint32_t shade(uint32_t col, int mode) {
		if (kSse2Shade) {
			// 16-bit lanes B,G,R,H: PosZ >>1; PosX/PosY/NegX (>>1)+(>>2); health<100 >>1; pack; filled bit
			const bool weak = (col >> 24) < 100u;
			const uint32_t half = (col >> 1) & 0x7f7f7f7fu, quarter = (col >> 2) & 0x3f3f3f3fu;
			uint32_t v = mode == 2 ? half : (mode == 1 ? half + quarter : col);
			v = weak ? (v >> 1) & 0x7f7f7f7fu : v;
			return __byte_perm(v, 0u, 0x3012) | (1u << 24); // H B G R -> H R G B
		} else {
			uint32_t c = __byte_perm(col, 0u, 0x4012); // 0 R G B
			c = mode == 2 ? (c & 0xfcfcfcu) >> 2 : (mode == 1 ? (c & 0xfefefeu) >> 1 : c);
			return c | (1u << 24);
		}
	}

	constexpr int kRcpSmemBits = 12; // rcpss tables up to 4096 entries (16 KB) are copied into shared memory per CTA

	template <bool kSse> struct LineCtx {
		osbn::Num<kSse> num;
		float cz, transScale, transOffset, vmax;
		uint32_t srcp; // shared-window address of the CTA's copy of the rcpss table, 0 = use the global one
		// reciprocal for the two hot uses (DDA distance, ToSpecialTan beyond the seam): same value as Num::rcpBig,
		// table entry fetched with ld.shared when the CTA holds a copy (keeps it out of the thrashed L1)
		__device__ __forceinline__ float rcpBig(float b) const {
#if OSB_RCP_SMEM
			if (kSse && srcp) {
				const uint32_t u = __float_as_uint(b), e = (u >> 23) & 0xff;
				uint32_t r;
				asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(srcp + (((u & 0x7fffff) >> (23 - num.t.rcpBits)) << 2)));
				if (e >= 253u)
					return osbn::rcpss(b, num.t);
				return __uint_as_float((u & 0x80000000u) | (r + ((127u - e) << 23)));
			}
#endif
			return num.rcpBig(b);
		}
		// transform lambda, SWMapRenderer.cpp:395-401 (ToSpecialTan :59-66 inlined).  Branch-free: the reciprocal is
		// always formed and selected by |v| > 1, so lanes on different sides of the seam share one instruction stream.
		__device__ __forceinline__ int transform(float invDist, int z) const {
			const float zv = __fsub_rn((float)z, cz);
			const float v = __fmul_rn(invDist, zv);
#if OSB_TRANSFORM_BRANCHY
			float t = v;
			if (fabsf(v) > 1.f)
				t = __fsub_rn(v < 0.f ? -2.f : 2.f, rcpBig(v));
#else
			const float r = rcpBig(v);
			float t = v;
			t = v < -1.f ? __fsub_rn(-2.f, r) : t;
			t = v > 1.f ? __fsub_rn(2.f, r) : t;
#endif
			float p = __fadd_rn(__fmul_rn(t, transScale), transOffset);
			// p < 0 ? 0 : p;  vmax < p ? vmax : p;  truncation.  p is never NaN here (finite * finite), and a NaN
			// would end as 0 either way (cvttss2si of the clamp result vs fmaxf dropping the NaN).
			p = fminf(fmaxf(p, 0.f), vmax);
			return __float2int_rz(p);
		}
	};

	// One thread = one yaw line (SWMapRenderer::BuildLine, SWMapRenderer.cpp:219-671).
	template <bool kSse, bool kSse2Shade>
#ifndef OSB_LINE_MINB
#define OSB_LINE_MINB 14
#endif
#ifndef OSB_STRIP_CLEAR
#define OSB_STRIP_CLEAR 0
#endif
#ifndef OSB_PAIR_STORE
#define OSB_PAIR_STORE 0
#endif
#ifndef OSB_MASK_FIRST
#define OSB_MASK_FIRST 1 // gather: test the written-pixel mask before loading the strip pixel (1: 1.68 ms) or load both at once (0: 1.91 ms)
#endif
	__global__ void __launch_bounds__(kLineThreads, OSB_LINE_MINB)
	  k_build_lines(const ViewParams *__restrict__ views, int nViews, const float2 *__restrict__ lineDirs,
	                const uint64_t *__restrict__ solid, const uint32_t *__restrict__ color,
	                osbn::SseTables tabs, uint2 *__restrict__ strips, int4 *__restrict__ lineParams,
	                uint32_t *__restrict__ lineMask, uint32_t totalLines, int lineRes, int lanesPerWarp) {
		extern __shared__ uint32_t s_filled[]; // (lineRes/32) words x kLineThreads
		// lanesPerWarp < 32 (small launches, e.g. one 1080p frame = a few thousand lines): the lines are spread over
		// more warps and SMs, and a warp's instruction stream is the union of fewer lanes' paths
		const uint32_t gl = (blockIdx.x * (kLineThreads / 32) + (threadIdx.x >> 5)) * (uint32_t)lanesPerWarp + (threadIdx.x & 31);
		const int words = lineRes >> 5;

		// LinePixel::Clear = {0, 10000.f} (:89-92).  The strips are NOT cleared in memory: a line publishes the bitmask
		// of the strip pixels it wrote (lineMask, lineRes bits per line) and the gather substitutes the clear value
		// for the others -- the clear was 3/4 of this kernel's DRAM writes.  OSB_STRIP_CLEAR=1 restores the clear.
		{
#if OSB_STRIP_CLEAR // (needs lanesPerWarp == 32)
			const uint32_t lineBase = blockIdx.x * kLineThreads;
			const uint32_t nl = min((uint32_t)kLineThreads, totalLines - lineBase);
			uint4 *dst = reinterpret_cast<uint4 *>(strips + (size_t)lineBase * lineRes);
			const uint32_t n4 = nl * (uint32_t)lineRes / 2;
			const uint4 clr = make_uint4(0u, __float_as_uint(10000.f), 0u, __float_as_uint(10000.f));
			for (uint32_t i = threadIdx.x; i < n4; i += kLineThreads)
				__stcs(dst + i, clr);
#endif
			for (int w = 0; w < words; w++)
				s_filled[w * kLineThreads + threadIdx.x] = 0;
		}
		// CTA copy of the rcpss table (SSE numerics, table small enough), behind the fill masks
		uint32_t srcp = 0;
#if OSB_RCP_SMEM
		if (kSse && tabs.rcpBits <= kRcpSmemBits) {
			uint32_t *s_rcp = s_filled + words * kLineThreads;
			for (int i = threadIdx.x; i < (1 << tabs.rcpBits); i += kLineThreads)
				s_rcp[i] = __ldg(tabs.rcp + i);
			srcp = (uint32_t)__cvta_generic_to_shared(s_rcp);
		}
#endif
		__syncthreads();
		if ((int)(threadIdx.x & 31) >= lanesPerWarp || gl >= totalLines)
			return;

		// which view does this line belong to (views are few per chunk: binary search on lineBase)
		int lo = 0, hi = nViews - 1;
		while (lo < hi) {
			const int mid = (lo + hi + 1) >> 1;
			if (views[mid].lineBase <= gl)
				lo = mid;
			else
				hi = mid - 1;
		}
		const ViewParams &V = views[lo];
		uint32_t *filled = s_filled + threadIdx.x;
		uint2 *strip = strips + (size_t)gl * lineRes;
		const float2 hd = lineDirs[gl];
		const float pi = 3.14159274101257324f; // static const float pi = M_PI

		// ---- pitch culling against the four side planes (:232-289) ----
		osbn::Num<kSse> num{tabs};
		float minPitch = -pi * 0.4999f, maxPitch = pi * 0.4999f;
#pragma unroll 1
		for (int k = 0; k < 4; k++) {
			const float px = V.frN[k][0], py = V.frN[k][1], pz = V.frN[k][2];
			if (px == 0.f && py == 0.f) {
				if (pz > 0.f)
					minPitch = minPitch < 0.f ? 0.f : minPitch;
				else
					maxPitch = 0.f < maxPitch ? 0.f : maxPitch;
			} else if (pz == 0.f) {
				// Vector3::Dot(plane, horz) with horz.z = 0: x*x' + y*y' + z*0
				const float d = __fadd_rn(__fadd_rn(__fmul_rn(px, hd.x), __fmul_rn(py, hd.y)), __fmul_rn(pz, 0.f));
				if (d < 0.f) {
					minPitch = 2.f;
					maxPitch = -2.f;
				}
			} else {
				// prj = (px,py,0).Normalize()
				float sc = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)), 0.f));
				if (sc != 0.f)
					sc = __fdiv_rn(1.f, sc);
				const float qx = __fmul_rn(px, sc), qy = __fmul_rn(py, sc), qz = __fmul_rn(0.f, sc);
				const float zv = fabsf(pz);
				const float cs = __fadd_rn(__fadd_rn(__fmul_rn(qx, hd.x), __fmul_rn(qy, hd.y)), __fmul_rn(qz, 0.f));
				const float cs2 = __fmul_rn(cs, cs);
				float ang = __fdiv_rn(__fmul_rn(__fmul_rn(zv, zv), __fsub_rn(1.f, cs2)), cs2);
				ang = __fmul_rn(-cs, num.sqrt(__fadd_rn(1.f, ang)));
				ang = __fdiv_rn(zv, ang);
				if (isnan(ang) || isinf(ang) || ang == 0.f)
					continue;
				ang = __fdiv_rn(num.sqrt(__fsub_rn(1.f, __fmul_rn(ang, ang))), ang);
				ang = osbm::atanf_glibc(ang);
				if (isnan(ang) || isinf(ang))
					continue;
				if (pz > 0.f) {
					const float c = __fsub_rn(ang, 0.01f);
					minPitch = minPitch < c ? c : minPitch;
				} else {
					const float c = __fadd_rn(-ang, 0.01f);
					maxPitch = c < maxPitch ? c : maxPitch;
				}
			}
		}
		// SpecialTan (:43-57)
		auto specialTan = [&](float v) -> float {
			if (v <= -pi * 0.5f)
				return -2.f;
			else if (v < -pi * 0.25f)
				return __fsub_rn(-2.f, __fdiv_rn(1.f, osbm::tanf_glibc(v)));
			else if (v < pi * 0.25f)
				return osbm::tanf_glibc(v);
			else if (v < pi * 0.5f)
				return __fsub_rn(2.f, __fdiv_rn(1.f, osbm::tanf_glibc(v)));
			return 2.f;
		};
		const float minTan = specialTan(minPitch);
		float maxTan = specialTan(maxPitch);
		{
			const float minDiff = __fdiv_rn((float)lineRes, 10000.f);
			if (maxTan < __fadd_rn(minTan, minDiff))
				maxTan = __fadd_rn(minTan, minDiff);
		}
		const float pitchScale = __fdiv_rn((float)lineRes, __fsub_rn(maxTan, minTan));
		lineParams[gl] = make_int4(__float_as_int(minTan), __float_as_int(pitchScale),
		                           osbn::x86_f2i(__fmul_rn(minTan, 65536.f)), osbn::x86_f2i(__fmul_rn(pitchScale, 65536.f)));
		uint32_t *maskOut = lineMask + (size_t)gl * words;
		if (minPitch >= maxPitch) { // culled: the strip stays cleared (:387-389)
			for (int w = 0; w < words; w++)
				maskOut[w] = 0;
			return;
		}

		// ---- DDA set-up (:309-351) ----
		float dirX = hd.x, dirY = hd.y;
		if (fabsf(dirY) < 1.e-4f)
			dirY = 1.e-4f;
		if (fabsf(dirX) < 1.e-4f)
			dirX = 1.e-4f;
		const float invDirX = __fdiv_rn(1.f, dirX), invDirY = __fdiv_rn(1.f, dirY);
		const bool posX = dirX > 0.f, posY = dirY > 0.f;
		const int invDirXI = abs(osbn::x86_f2i(__fmul_rn(invDirX, 256.f)));
		const int invDirYI = abs(osbn::x86_f2i(__fmul_rn(invDirY, 256.f)));
		const uint32_t dirXI = (uint32_t)abs(osbn::x86_f2i(__fmul_rn(dirX, 512.f)));
		const uint32_t dirYI = (uint32_t)abs(osbn::x86_f2i(__fmul_rn(dirY, 512.f)));

		const float cx = V.origin[0], cy = V.origin[1], cz = V.origin[2];
		const int icz = osbn::x86_f2i(floorf(cz));
		int rx = osbn::x86_f2i(__fmul_rn(cx, 512.f)), ry = osbn::x86_f2i(__fmul_rn(cy, 512.f));
		int irx = rx >> 9, iry = ry >> 9;

		const float fogDist = 128.f;
		float distance = 1.e-20f;
		float invDist = __fdiv_rn(1.f, distance);

		LineCtx<kSse> L;
		L.num = num;
		L.cz = cz;
		L.transScale = pitchScale; // same expression as line.pitchScale (:303,358)
		L.transOffset = __fmul_rn(-minTan, L.transScale);
		L.vmax = __fadd_rn((float)lineRes, 0.5f);
		L.srcp = srcp;

		// zscale = Dot(horizonDir, viewAxis[2]) (horizonDir.z = 0), heightScale, depthBias (:403-414)
		const float zscale = __fadd_rn(__fadd_rn(__fmul_rn(hd.x, V.front[0]), __fmul_rn(hd.y, V.front[1])),
		                               __fmul_rn(0.f, V.front[2]));
		const float heightScale = V.front[2];
		const float depthBias = __fmul_rn(-cz, heightScale);

		// z > icz / z < icz as bit masks
		const uint64_t belowEye = icz >= 63 ? 0ull : (icz < 0 ? ~0ull : (~0ull << (icz + 1)));
		const uint64_t aboveEye = icz <= 0 ? 0ull : (icz >= 64 ? ~0ull : ((1ull << icz) - 1ull));

		uint32_t tOld = tiled_index(irx & 511, iry & 511);
		uint64_t sOld = solid[tOld];
		int count = 1;
		uint32_t cnt2 = 1024; // static_cast<int>(fogDist * 8.f)

		// One DDA step (:426-522) depends only on (rx, ry), never on map data, so the step after the current
		// one is computed ahead and its column mask is requested before the current step's spans are
		// processed: the dependent 8-byte load is off the critical path.  Both crossing cases are formed with
		// selects so that lanes crossing x and lanes crossing y stay in one instruction stream.
		struct Step {
			int rx, ry, irx, iry, wallMode;
			float distance;
			uint32_t t;
		};
		const int sx = posX ? 1 : -1, sy = posY ? 1 : -1;
		const int edgeX = posX ? 0 : 1, edgeY = posY ? 0 : 1;
		// wall face of an x crossing is NegX / PosX (both 3/4-shaded, mode 1); of a y crossing NegY (mode 0) / PosY (1)
		const int wallModeY = posY ? 0 : 1;
		auto advance = [&](int rx0, int ry0, int irx0, int iry0, float dist0) -> Step {
			Step n;
			const uint32_t tx = (uint32_t)((posX ? 512 - (rx0 & 511) : (rx0 & 511)) * invDirXI);
			const uint32_t ty = (uint32_t)((posY ? 512 - (ry0 & 511) : (ry0 & 511)) * invDirYI);
			const bool xs = tx < ty;
			const uint32_t tt = xs ? tx : ty;
			const int d = (int)(((xs ? dirYI : dirXI) * tt) >> 17);
			n.irx = irx0 + (xs ? sx : 0);
			n.iry = iry0 + (xs ? 0 : sy);
			n.rx = xs ? (n.irx << 9) - edgeX : rx0 + sx * d;
			n.ry = xs ? ry0 + sy * d : (n.iry << 9) - edgeY;
			n.distance = __fadd_rn(dist0, __fmul_rn((float)tt, 1.f / 512.f / 256.f));
			n.wallMode = xs ? 1 : wallModeY;
			n.t = tiled_index(n.irx & 511, n.iry & 511);
			return n;
		};
		Step nx = advance(rx, ry, irx, iry, distance);
		uint64_t sNext = solid[nx.t];

		while (distance < fogDist && (--cnt2) > 0) {
			// ---- take the precomputed step, start the one after it ----
			const Step cur = nx;
			const uint64_t sNew = sNext;
			STAT(0, 1);
			distance = cur.distance;
			nx = advance(cur.rx, cur.ry, cur.irx, cur.iry, cur.distance);
			sNext = solid[nx.t];
			const uint32_t tNew = cur.t;

			const float oldInvDist = invDist;
			invDist = L.rcpBig(distance); // distance is in [1e-20, ~200]
			const float medDist = __fadd_rn(__fmul_rn(distance, zscale), depthBias);

			// ---- spans of this step, one merged work list per lane, in the reference's order:
			//   floor tops of the cell being left (ascending z, :584-603), its ceilings (descending z, :604-622),
			//   then the walls of the cell being entered for the crossed face (ascending z, :627-657).
			// One loop body serves all three kinds (ceilings are kept bit-reversed so that "next" is always the
			// lowest set bit), and everything up to the landing test is straight-line code.
			uint64_t tops = sOld & ~(sOld << 1) & belowEye;                                // solid, air above, below the eye
			uint64_t botsR = __brevll(sOld & ~(sOld >> 1) & ~(1ull << 63) & aboveEye);      // solid, air below, above the eye
			uint64_t walls = sNew & ~sOld;                                                  // solid there, air in the cell left
			while (tops | botsR | walls) {
				const bool isTop = tops != 0, isBot = !isTop && botsR != 0;
				const uint64_t src = isTop ? tops : (isBot ? botsR : walls);
				const int bit = __ffsll((long long)src) - 1;
				const uint64_t rest = src & (src - 1);
				tops = isTop ? rest : tops;
				botsR = isBot ? rest : botsR;
				walls = (isTop || isBot) ? walls : rest;
				const int z = isBot ? 63 - bit : bit;
				const bool isWall = !(isTop || isBot);
#if OSB_COLOR_EARLY
				// colour is requested before the projection math; it is only used when the span lands
				const uint32_t rawCol = color[(size_t)(isWall ? tNew : tOld) * 64 + z];
#endif
				const int zA = isBot ? z + 1 : z;      // first transform: (invDist, zA)
				const int zB = isTop ? z : z + 1;      // second: (oldInvDist or invDist, zB)
				const int pA = L.transform(invDist, zA);
				const int pB = L.transform(isWall ? invDist : oldInvDist, zB);
				const int from = isBot ? pB : pA, to = isBot ? pA : pB;
				if (from < to) {
					const int w0 = from >> 5, w1 = (to - 1) >> 5;
					const uint32_t mFirst = 0xffffffffu << (from & 31), mLast = 0xffffffffu >> (31 - ((to - 1) & 31));
					uint2 px;
					bool havePx = false;
					for (int w = w0; w <= w1; w++) {
						uint32_t m = 0xffffffffu;
						m = w == w0 ? m & mFirst : m;
						m = w == w1 ? m & mLast : m;
						const uint32_t curMask = filled[w * kLineThreads];
						uint32_t todo = m & ~curMask;
						if (!todo)
							continue;
						filled[w * kLineThreads] = curMask | todo;
						if (!havePx) {
#if OSB_EXP == 1
							const uint32_t rawCol = 0x64000000u | (uint32_t)z * 0x030201u;
#elif !OSB_COLOR_EARLY
							const uint32_t rawCol = color[(size_t)(isWall ? tNew : tOld) * 64 + z];
#endif
							const int mode = isTop ? 0 : (isBot ? 2 : cur.wallMode);
							px = make_uint2(shade<kSse2Shade>(rawCol, mode),
							                __float_as_uint(__fadd_rn(medDist, __fmul_rn((float)zA, heightScale))));
							havePx = true;
						}
						uint2 *dst = strip + (w << 5);
#if OSB_PAIR_STORE
						// two adjacent strip pixels (even-aligned) go out as one 16-byte store
						uint32_t pairs = todo & (todo >> 1) & 0x55555555u;
						todo &= ~(pairs | (pairs << 1));
						const uint4 px2 = make_uint4(px.x, px.y, px.x, px.y);
						while (pairs) {
							const int b = __ffs(pairs) - 1;
							pairs &= pairs - 1;
							__stcs(reinterpret_cast<uint4 *>(dst + b), px2);
						}
						while (todo) {
							const int b = __ffs(todo) - 1;
							todo &= todo - 1;
							__stcs(dst + b, px);
						}
#else
						do {
							const int b = __ffs(todo) - 1;
							todo &= todo - 1;
#if OSB_EXP == 2
							if (px.x == 0xdeadbeefu)
#endif
							__stcs(dst + b, px);
						} while (todo);
#endif
					}
				}
			}
			sOld = sNew;
			tOld = tNew;
			// ---- pitch cull every 4th step (:661-667) ----
			if ((--count) == 0) {
				if ((L.transform(invDist, 0) >= lineRes - 1 && icz >= 0) || L.transform(invDist, 63) <= 0)
					break;
				count = 4;
			}
		}
		for (int w = 0; w < words; w++)
			maskOut[w] = filled[w * kLineThreads];
	}

	// ---- fastATan / fastATan2 (SWMapRenderer.cpp:697-742); tab = sm | lg | smN | lgN, 5000 each ----
	template <bool kSse>
	__device__ __forceinline__ uint32_t fast_atan(const osbn::Num<kSse> &num, const uint16_t *__restrict__ tab, float v) {
		int idx, base;
		if (v < 0.f) {
			if (v > -1.f) {
				idx = osbn::x86_f2i(__fmul_rn(v, -4096.f));
				base = 10000;
			} else {
				idx = osbn::x86_f2i(num.div(-4096.f, v));
				base = 15000;
			}
		} else {
			if (v < 1.f) {
				idx = osbn::x86_f2i(__fmul_rn(v, 4096.f));
				base = 0;
			} else {
				idx = osbn::x86_f2i(num.div(4096.f, v));
				base = 5000;
			}
		}
		idx = min(max(idx, 0), 4999); // memory safety only; in range for every finite input
		return __ldg(tab + base + idx);
	}
	template <bool kSse>
	__device__ __forceinline__ uint32_t fast_atan2(const osbn::Num<kSse> &num, const uint16_t *__restrict__ tab, float y,
	                                               float x) {
		if (x == 0.f)
			return y > 0.f ? 16384u : 49152u; // (uint16_t)-16384
		else if (x > 0.f)
			return fast_atan(num, tab, num.div(y, x));
		return (fast_atan(num, tab, num.div(y, x)) + 32768u) & 0xffffu;
	}

	constexpr int kGatherWarps = 4;
	constexpr int kSegRows = 8; // 8x8 blocks per warp task (vertical)

	// truncating (C) division by 2^sh without the 40-instruction generic idiv
	__device__ __forceinline__ int tdiv_pow2(int x, int sh) { return (x + ((x >> 31) & ((1 << sh) - 1))) >> sh; }

	// ApplyFog blend of one pixel (SWRenderer.cpp:381-407 scalar, :462-510 SSE2).  col = B | G<<8 | R<<16 | X<<24.
	// B and R share one multiply: each 8-bit channel times a factor <= 256 stays below 2^16.
	template <bool kSse2Fog>
	__device__ __forceinline__ uint32_t fog_blend(uint32_t col, float dist, uint32_t fogBR, uint32_t fogG8) {
		int factor;
		if (kSse2Fog) {
			float d = dist > 0.f ? dist : 0.f; // maxps / minps return the 2nd operand on NaN
			d = d < 256.f ? d : 256.f;
			factor = osbn::x86_f2i_rn(d);
		} else {
			factor = min(osbn::x86_f2i(dist), 256);
			factor = max(0, factor);
		}
		const uint32_t f1 = (uint32_t)factor, f2 = (uint32_t)(256 - factor);
		const uint32_t br = (((col & 0x00ff00ffu) * f2 + fogBR * f1) >> 8) & 0x00ff00ffu;
		const uint32_t gg = (((col & 0x0000ff00u) * f2 + fogG8 * f1) >> 8) & 0x0000ff00u;
		uint32_t out = br | gg;
		if (kSse2Fog) // lane 3 of the SSE2 path keeps (X * f2) >> 8
			out |= (((col >> 24) * f2) << 16) & 0xff000000u;
		return out;
	}

	// RenderFinal + ApplyFog.  One warp per (view, 32-pixel column group = four 8x8 block columns, segment of
	// kSegRows block rows); one lane per pixel column, so every row is written as one 128-byte run per buffer.
	// Per block row the 16 block corners (4 blocks x 4) are evaluated by 16 lanes in one instruction stream
	// and shared by shuffle; per pixel only the reference's integer interpolation, the gather and the blend remain.
	template <bool kSse, bool kSse2Fog>
	// (a min-blocks bound was measured and is not used: 10 / 12 / 16 CTAs per SM spill, 1.85 / 1.94 / 2.04 ms; and an explicit
	// bound of 1 lets ptxas take more registers than the plain form below: 1.90 vs 1.68 ms)
	__global__ void __launch_bounds__(kGatherWarps * 32)
	  k_gather_fog(const ViewParams *__restrict__ views, int nViews, FrameGeom g, const float *__restrict__ tabs,
	               const uint64_t *__restrict__ solid, const uint16_t *__restrict__ atanTab, osbn::SseTables sse,
	               const uint2 *__restrict__ strips, const int4 *__restrict__ lineParams,
	               const uint32_t *__restrict__ lineMask, int lineRes, uint32_t *__restrict__ fb,
	               float *__restrict__ depthBuf) {
		__shared__ float s_fog[kGatherWarps][kSegRows * 2][8];
		const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
		const int colBlocks = g.fw >> 3, rowBlocks = g.fh >> 3;
		const int colGroups = (colBlocks + 3) >> 2;
		const int segs = (rowBlocks + kSegRows - 1) / kSegRows;
		const uint32_t task = blockIdx.x * kGatherWarps + wid;
		const uint32_t perView = (uint32_t)colGroups * segs;
		if (task >= perView * (uint32_t)nViews)
			return;
		const int vi = task / perView;
		const uint32_t rem = task - (uint32_t)vi * perView;
		const int seg = rem / colGroups, cg = rem - seg * colGroups;
		const ViewParams &V = views[vi];
		osbn::Num<kSse> num{sse};
		const float *vxTab = tabs + V.tabBase, *vyTab = vxTab + (g.fw >> 2), *v1Tab = vyTab + (g.fh >> 2);
		uint32_t *fbV = fb + (size_t)vi * g.fw * g.fh;
		float *dbV = depthBuf + (size_t)vi * g.fw * g.fh;
		const int by0 = seg * kSegRows, by1 = min(by0 + kSegRows, rowBlocks);

		const int bx = (cg << 2) + (lane >> 3);  // this lane's block column
		const bool colValid = bx < colBlocks;    // a partial last group when fw is not a multiple of 32
		const int bxc = colValid ? bx : colBlocks - 1;
		const int px = lane & 7;                 // pixel column inside the block
		const int x = (bxc << 3) + px;

		// SWMapRenderer::Render :1181-1184: eye inside a solid voxel (or no map at all) -> terrain skipped, fb keeps
		// the clear colour and the depth buffer keeps its previous contents
		bool skip;
		{
			const int ex = osbn::x86_f2i(floorf(V.origin[0])), ey = osbn::x86_f2i(floorf(V.origin[1])),
			          ez = osbn::x86_f2i(floorf(V.origin[2]));
			skip = g.noMap || ez >= 64 || (ez >= 0 && ((solid[tiled_index(ex & 511, ey & 511)] >> ez) & 1ull));
		}

		// ---- fog scale of every 4x4 pixel block this task covers (SWRenderer.cpp:372-380) ----
		{
			const int nRows4 = (by1 - by0) * 2;
			for (int i = lane; i < nRows4 * 8; i += 32) {
				const int r4 = i >> 3, c4 = i & 7;
				const int cx4 = min((cg << 3) + c4, (g.fw >> 2) - 1);
				const float vx = vxTab[cx4], vy = vyTab[(by0 << 1) + r4];
				float ds = __fadd_rn(__fadd_rn(1.f, __fmul_rn(vx, vx)), __fmul_rn(vy, vy));
				ds = __fmul_rn(ds, __fmul_rn(num.rsqrt(ds), V.fogScale));
				s_fog[wid][r4][c4] = ds;
			}
			__syncwarp();
		}
		const uint32_t fogBR = g.fogB | (g.fogR << 16), fogG8 = g.fogG << 8;

		// per-view deltas (:757-775)
		const int under = g.under, hSh = under == 1 ? 3 : (under == 2 ? 2 : 1); // log2(8 / under)
		const int pxq = px - (px % under), xs = pxq / under; // undersampling: pixels share their group leader's sample
		const float sD = __fdiv_rn(__fmul_rn(V.fovYt, 2.f), (float)g.fh);
		const float sR = __fmul_rn(__fdiv_rn(__fmul_rn(V.fovXt, 2.f), (float)g.fw), (float)under);
		float dD[3], dR[3], dDL[3], dRL[3], v2[3];
#pragma unroll
		for (int k = 0; k < 3; k++) {
			dD[k] = __fmul_rn(-V.up[k], sD);
			dR[k] = __fmul_rn(V.right[k], sR);
			dDL[k] = __fmul_rn(dD[k], 8.f);
			dRL[k] = __fmul_rn(dR[k], (float)(8 / under));
			v2[k] = v1Tab[bxc * 3 + k];
		}
		for (int i = 0; i < by0; i++) { // replay v2 += deltaDownLarge (:1037)
#pragma unroll
			for (int k = 0; k < 3; k++)
				v2[k] = __fadd_rn(v2[k], dDL[k]);
		}
		const uint32_t numLines = (uint32_t)V.numLines;
		const uint32_t yawScale2 = (uint32_t)V.yawScale2;
		const int yawMin2 = V.yawMin2;
		const int lineShift = 31 - __clz(lineRes);
		const uint2 *stripsV = strips + (size_t)V.lineBase * lineRes;
		const uint32_t *maskV = lineMask + (size_t)V.lineBase * (lineRes >> 5); // written strip pixels, bit per pixel
		const uint2 kClearPix = make_uint2(0u, __float_as_uint(10000.f));       // LinePixel::Clear
		const int4 *lpV = lineParams + V.lineBase;
		const int blkLane = lane & ~7;

		for (int by = by0; by < by1; by++) {
			size_t o = (size_t)(by * 8) * g.fw + x;
			const float *fogRow = &s_fog[wid][(by - by0) * 2][lane >> 2];
			if (skip) {
				if (colValid) {
#pragma unroll
					for (int py = 0; py < 8; py++, o += g.fw) {
						const float dep = dbV[o];
						fbV[o] = g.noFog ? 0x7f7f7fu : fog_blend<kSse2Fog>(0x7f7f7fu, __fmul_rn(dep, fogRow[(py >> 2) * 8]), fogBR, fogG8);
					}
				}
				continue;
			}
			const bool slow = v2[2] > 0.99f || v2[2] < -0.99f; // uniform within a block (8 lanes)
			// ---- block corners: lanes 0..3 of every block (:810-835) ----
			int yaw = 0, pit = 0;
			if (px < 4 && !slow) {
				const int cr = px & 1, cd = (px >> 1) & 1;
				float c[3];
#pragma unroll
				for (int k = 0; k < 3; k++) {
					c[k] = v2[k];
					if (cr)
						c[k] = __fadd_rn(c[k], dRL[k]);
					if (cd)
						c[k] = __fadd_rn(c[k], dDL[k]);
				}
				yaw = (int)((((int)fast_atan2(num, atanTab, c[1], c[0]) - yawMin2) & 0xffff) << 8);
				float pitch = __fmul_rn(c[2], num.rsqrt(__fadd_rn(__fmul_rn(c[0], c[0]), __fmul_rn(c[1], c[1]))));
				pitch = num.toSpecialTan(pitch);
				pit = osbn::x86_f2i(__fmul_rn(pitch, 65536.f * 8192.f));
			}
			const int yaw1 = __shfl_sync(0xffffffffu, yaw, blkLane), yaw2 = __shfl_sync(0xffffffffu, yaw, blkLane + 1);
			const int yaw3 = __shfl_sync(0xffffffffu, yaw, blkLane + 2), yaw4 = __shfl_sync(0xffffffffu, yaw, blkLane + 3);
			const int pit1 = __shfl_sync(0xffffffffu, pit, blkLane), pit2 = __shfl_sync(0xffffffffu, pit, blkLane + 1);
			const int pit3 = __shfl_sync(0xffffffffu, pit, blkLane + 2), pit4 = __shfl_sync(0xffffffffu, pit, blkLane + 3);
			if (colValid) {
				if (!slow) {
					// bilinear interpolation in integers; `<<8>>8` is the reference's phase unwrap (:837-881)
					const int yawDiff1 = tdiv_pow2(((yaw2 - yaw1) << 8) >> 8, hSh), yawDiff2 = tdiv_pow2(((yaw4 - yaw3) << 8) >> 8, hSh);
					const int pitDiff1 = tdiv_pow2(pit2 - pit1, hSh), pitDiff2 = tdiv_pow2(pit4 - pit3, hSh);
					const int yawA = yaw1 + xs * yawDiff1, yawB = yaw3 + xs * yawDiff2;
					const int pitA = pit1 + xs * pitDiff1, pitB = pit3 + xs * pitDiff2;
					const int dYaw = tdiv_pow2(((yawB - yawA) << 8) >> 8, 3), dPit = tdiv_pow2(pitB - pitA, 3);
					int yawC = yawA, pitC = pitA;
#pragma unroll
					for (int py = 0; py < 8; py++, o += g.fw, yawC += dYaw, pitC += dPit) {
						uint32_t yawIndex = (uint32_t)((yawC << 8) >> 16);
						yawIndex = (yawIndex * yawScale2) >> 16;
						yawIndex = (yawIndex * numLines) >> 16;
						yawIndex = min(yawIndex, numLines - 1); // memory safety; the reference would read a stale line
						const int4 lp = __ldg(lpV + yawIndex);
						int pitchIndex = (pitC >> 13) - lp.z;
						pitchIndex = __mulhi(pitchIndex, lp.w);
						pitchIndex &= lineRes - 1;
#if OSB_STRIP_CLEAR
						const uint2 pix = __ldg(stripsV + (((size_t)yawIndex << lineShift) + pitchIndex));
#elif OSB_MASK_FIRST
						const uint32_t mw = __ldg(maskV + (((size_t)yawIndex << (lineShift - 5)) + (pitchIndex >> 5)));
						uint2 pix = kClearPix;
						if ((mw >> (pitchIndex & 31)) & 1u)
							pix = __ldg(stripsV + (((size_t)yawIndex << lineShift) + pitchIndex));
#else
						uint2 pix = __ldg(stripsV + (((size_t)yawIndex << lineShift) + pitchIndex));
						{
							const uint32_t mw = __ldg(maskV + (((size_t)yawIndex << (lineShift - 5)) + (pitchIndex >> 5)));
							pix = ((mw >> (pitchIndex & 31)) & 1u) ? pix : kClearPix;
						}
#endif
						const float dep = __uint_as_float(pix.y);
						__stcs(fbV + o, g.noFog ? pix.x : fog_blend<kSse2Fog>(pix.x, __fmul_rn(dep, fogRow[(py >> 2) * 8]), fogBR, fogG8));
						__stcs(dbV + o, dep);
					}
				} else {
					// exact per-pixel path near the poles (:940-1033)
					float v4[3];
#pragma unroll
					for (int k = 0; k < 3; k++)
						v4[k] = v2[k];
					for (int i = 0; i < xs; i++) {
#pragma unroll
						for (int k = 0; k < 3; k++)
							v4[k] = __fadd_rn(v4[k], dR[k]);
					}
#pragma unroll 1
					for (int py = 0; py < 8; py++, o += g.fw) {
						uint32_t yawIndex = (uint32_t)(((int)fast_atan2(num, atanTab, v4[1], v4[0]) - yawMin2) & 0xffff);
						yawIndex = (yawIndex * yawScale2) >> 16;
						yawIndex = (yawIndex * numLines) >> 16;
						yawIndex = min(yawIndex, numLines - 1);
						const int4 lp = __ldg(lpV + yawIndex);
						float pitch = __fmul_rn(v4[2], num.rsqrt(__fadd_rn(__fmul_rn(v4[0], v4[0]), __fmul_rn(v4[1], v4[1]))));
						pitch = num.toSpecialTan(pitch);
						pitch = __fmul_rn(__fsub_rn(pitch, __int_as_float(lp.x)), __int_as_float(lp.y));
						const int pitchIndex = osbn::x86_f2i(pitch) & (lineRes - 1);
#if OSB_STRIP_CLEAR
						const uint2 pix = __ldg(stripsV + (((size_t)yawIndex << lineShift) + pitchIndex));
#elif OSB_MASK_FIRST
						const uint32_t mw = __ldg(maskV + (((size_t)yawIndex << (lineShift - 5)) + (pitchIndex >> 5)));
						uint2 pix = kClearPix;
						if ((mw >> (pitchIndex & 31)) & 1u)
							pix = __ldg(stripsV + (((size_t)yawIndex << lineShift) + pitchIndex));
#else
						uint2 pix = __ldg(stripsV + (((size_t)yawIndex << lineShift) + pitchIndex));
						{
							const uint32_t mw = __ldg(maskV + (((size_t)yawIndex << (lineShift - 5)) + (pitchIndex >> 5)));
							pix = ((mw >> (pitchIndex & 31)) & 1u) ? pix : kClearPix;
						}
#endif
						const float dep = __uint_as_float(pix.y);
						__stcs(fbV + o, g.noFog ? pix.x : fog_blend<kSse2Fog>(pix.x, __fmul_rn(dep, fogRow[(py >> 2) * 8]), fogBR, fogG8));
						__stcs(dbV + o, dep);
#pragma unroll
						for (int k = 0; k < 3; k++)
							v4[k] = __fadd_rn(v4[k], dD[k]);
					}
				}
			}
#pragma unroll
			for (int k = 0; k < 3; k++)
				v2[k] = __fadd_rn(v2[k], dDL[k]);
		}
	}

	// Device-side recurrences for large batches: one warp per view, one lane per role.
	__global__ void k_view_prep(const ViewParams *__restrict__ views, int nViews, FrameGeom g, float2 *lineDirs,
	                            float *tabs) {
		const int vi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
		const int lane = threadIdx.x & 31;
		if (vi >= nViews)
			return;
		const ViewParams &V = views[vi];
		float *vx = tabs + V.tabBase, *vy = vx + (g.fw >> 2), *v1 = vy + (g.fh >> 2);
		if (lane == 0)
			prep_line_dirs(V, lineDirs + V.lineBase);
		else if (lane == 1)
			prep_fog_vx(V, g, vx);
		// the emulated SW threads (<= 32) are spread over the lanes for the vy bands and v1 columns
		for (int t = lane; t < g.swThreads; t += 32)
			prep_fog_vy(V, g, t, vy);
		for (int t = 31 - lane; t < g.swThreads; t += 32)
			prep_v1(V, g, t, v1);
	}

	static int terrain_init(osb200_ctx *c) {
		if (c->terrain)
			return 0;
		auto *T = new TerrainState();
		c->terrain = T;
		// AtanTable (SWMapRenderer.cpp:673-696), built with the host libm exactly like the reference
		std::vector<uint16_t> tab(20000);
		auto toFixed = [](float v) -> uint16_t {
			v /= (M_PI * 2.f);
			v *= 65536.f;
			int i = static_cast<int>(v);
			return static_cast<uint16_t>(i & 65535);
		};
		for (int i = 0; i < 5000; i++) {
			tab[i] = toFixed(atanf(i / 4096.f));
			tab[5000 + i] = toFixed(atanf(1.f / ((i + .5f) / 4096.f)));
			tab[10000 + i] = toFixed(-atanf(i / 4096.f));
			tab[15000 + i] = toFixed(-atanf(1.f / ((i + .5f) / 4096.f)));
		}
		OSB_CUDA(cudaMalloc(&T->d_atan, tab.size() * 2));
		OSB_CUDA(cudaMemcpy(T->d_atan, tab.data(), tab.size() * 2, cudaMemcpyHostToDevice));
		OSB_CUDA(cudaEventCreateWithFlags(&T->stageFree, cudaEventDisableTiming));
		OSB_CUDA(cudaStreamCreateWithFlags(&T->copyStream, cudaStreamNonBlocking));
		OSB_CUDA(cudaEventCreateWithFlags(&T->fork, cudaEventDisableTiming));
		OSB_CUDA(cudaEventCreateWithFlags(&T->copyDone, cudaEventDisableTiming));
		for (ChunkBufs &b : T->set) {
			OSB_CUDA(cudaStreamCreateWithFlags(&b.stream, cudaStreamNonBlocking));
			OSB_CUDA(cudaEventCreateWithFlags(&b.done, cudaEventDisableTiming));
		}
		return 0;
	}

	void terrain_destroy(osb200_ctx *c) {
		TerrainState *T = c->terrain;
		if (!T)
			return;
		cudaFree(T->d_views);
		cudaFreeHost(T->h_views);
		for (ChunkBufs &b : T->set) {
			cudaFree(b.d_lineDirs);
			cudaFree(b.d_lineParams);
			cudaFree(b.d_strips);
			cudaFree(b.d_lineMask);
			cudaFree(b.d_tabs);
			if (b.done)
				cudaEventDestroy(b.done);
			if (b.stream)
				cudaStreamDestroy(b.stream);
		}
		cudaFreeHost(T->h_tabs);
		cudaFreeHost(T->h_lineDirs);
		cudaFree(T->d_atan);
		cudaFree(T->d_fb);
		cudaFree(T->d_depth);
		if (T->stageFree)
			cudaEventDestroy(T->stageFree);
		if (T->fork)
			cudaEventDestroy(T->fork);
		if (T->copyDone)
			cudaEventDestroy(T->copyDone);
		if (T->copyStream)
			cudaStreamDestroy(T->copyStream);
		delete T;
		c->terrain = nullptr;
	}

	// Host half of RenderInner (SWMapRenderer.cpp:1057-1133) + BuildFrustrum (SWRenderer.cpp:207-227) +
	// ApplyFog's scalars (SWRenderer.cpp:345-357) for one view, with the host libm.
	static void fill_view(const osb200_scene &s, int fw, int fh, int under, float fogDist, ViewParams &v) {
		static const float pi = M_PI;
		std::memset(&v, 0, sizeof(v));
		for (int k = 0; k < 3; k++) {
			v.origin[k] = s.origin[k];
			v.right[k] = s.axis[0][k];
			v.up[k] = s.axis[1][k];
			v.front[k] = s.axis[2][k];
		}
		const float xCos = cosf(s.fovX * .5f), xSin = sinf(s.fovX * .5f);
		const float yCos = cosf(s.fovY * .5f), ySin = sinf(s.fovY * .5f);
		for (int k = 0; k < 3; k++) {
			v.frN[0][k] = v.front[k] * xSin - v.right[k] * xCos;
			v.frN[1][k] = v.front[k] * xSin + v.right[k] * xCos;
			v.frN[2][k] = v.front[k] * ySin - v.up[k] * yCos;
			v.frN[3][k] = v.front[k] * ySin + v.up[k] * yCos;
		}
		const float fovX = tanf(s.fovX * 0.5f), fovY = tanf(s.fovY * 0.5f);
		v.fovXt = fovX;
		v.fovYt = fovY;
		const float fovDiag = sqrtf(fovX * fovX + fovY * fovY);
		const float fovDiagAng = atanf(fovDiag);
		const float pitch = asinf(v.front[2]);
		float yawMin, yawMax;
		if (fabsf(pitch) >= pi * 0.49f - fovDiagAng) { // pole is visible
			yawMin = 0.f;
			yawMax = pi * 2.f;
		} else {
			const float yaw = atan2l(v.front[1], v.front[0]);
			yawMin = yaw - pi * .5f;
			yawMax = yaw + pi * .5f;
		}
		float interval = static_cast<float>(fh);
		interval = fovY * 2.f / interval;
		int lineRes = fh;
		for (int i = lineRes, j = 1; j <= i; j <<= 1)
			lineRes = j;
		size_t numLines = static_cast<size_t>((yawMax - yawMin) / interval);
		numLines /= under;
		if (numLines < 8)
			numLines = 8;
		if (numLines > 65536)
			numLines = 65536;
		v.yawMin = yawMin;
		v.yawMax = yawMax;
		v.numLines = (int)numLines;
		v.lineRes = lineRes;
		const float scl = (yawMax - yawMin) / numLines;
		v.cosYawMin = cosf(yawMin);
		v.sinYawMin = sinf(yawMin);
		v.cosScl = cosf(scl);
		v.sinScl = sinf(scl);
		const float yawScale = 65536.f / (pi * 2.f);
		v.yawScale2 = static_cast<int32_t>(pi * 2.f / (yawMax - yawMin) * 65536.f);
		v.yawMin2 = static_cast<int32_t>(yawMin * yawScale);
		v.fogScale = 255.f / fogDist;
		v.dvx = -fovX * 2.f / static_cast<float>(fw / 4);
		v.dvy = -fovY * 2.f / static_cast<float>(fh / 4);
	}
} // namespace osb

namespace osb {
	template <class T> static int grow(T *&p, size_t &cap, size_t want, bool pinned = false) {
		if (want <= cap)
			return 0;
		if (p) {
			if (pinned)
				cudaFreeHost(p);
			else
				cudaFree(p);
		}
		p = nullptr;
		cap = 0;
		size_t n = want + (want >> 2);
		if (pinned)
			OSB_CUDA(cudaMallocHost((void **)&p, n * sizeof(T)));
		else
			OSB_CUDA(cudaMalloc((void **)&p, n * sizeof(T)));
		cap = n;
		return 0;
	}

	static size_t env_mb(const char *name, long dflt) {
		const char *e = getenv(name);
		long mb = e ? atol(e) : 0;
		return (size_t)(mb > 0 ? mb : dflt) << 20;
	}
	// Strip bytes per chunk.  Measured on B200: every launch ends in a latency-bound tail, so chunks must be large
	// for the tail to be small against the launch, and there must be several of them for the two streams to overlap
	// one chunk's tail with the other chunk's kernels.
	static size_t strip_budget_bytes() {
		static size_t v = env_mb("OSB200_STRIP_MB", 3072);
		return v;
	}
	static bool dual_stream() {
		static int v = -1;
		if (v < 0) {
			const char *e = getenv("OSB200_DUAL_STREAM");
			v = e ? atoi(e) != 0 : 1;
		}
		return v != 0;
	}

	// host_fb / host_depth (nullable): when given, each chunk's finished frames are copied to the host on a second
	// stream while the next chunk renders (PCIe overlaps the kernels); the caller synchronises copyDone.
	int render_impl(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
	                float fogDist, uint32_t *d_fb, float *d_depth, uint32_t *host_fb, float *host_depth, bool deferFog) {
		if (!c) {
			set_error("osb200_render: no context");
			return -1;
		}
		if (nViews == 0)
			return 0;
		if (!scenes || !fogColor || !d_fb || !d_depth || nViews < 0) {
			set_error("osb200_render: null argument");
			return -1;
		}
		if (w <= 0 || h <= 0 || (w & 7) || (h & 7)) {
			set_error("Framebuffer size is not multiple of 8."); // SWRenderer.cpp:105-107
			return -1;
		}
		if (h >= 8192 || w > 32768) {
			set_error("osb200_render: framebuffer too large (h < 8192, w <= 32768)");
			return -1;
		}
		OSB_CUDA(cudaSetDevice(c->device));
		if (terrain_init(c))
			return -2;
		TerrainState *T = c->terrain;
		const int under = c->opts.undersampling;
		FrameGeom g;
		g.fw = w;
		g.fh = h;
		g.under = under;
		g.swThreads = c->opts.sw_threads;
		g.sse2Fog = g.sse2Shade = c->opts.numerics == OSB200_NUM_SSE2;
		g.noMap = c->mapLoaded ? 0 : 1;
		g.noFog = deferFog ? 1 : 0;
		auto toFixed8 = [](float v) { // SWUtils.h:68-71
			int i = static_cast<int>(v * 255.f + .5f);
			return (uint32_t)std::max(std::min(i, 255), 0);
		};
		g.fogR = toFixed8(fogColor[0]);
		g.fogG = toFixed8(fogColor[1]);
		g.fogB = toFixed8(fogColor[2]);

		// the pinned staging buffers may still be in flight from the previous call
		OSB_CUDA(cudaEventSynchronize(T->stageFree));
		if (grow(T->h_views, T->viewCap, (size_t)nViews, true))
			return -2;
		if (grow(T->d_views, T->dViewCap, (size_t)nViews))
			return -2;
		const uint32_t tabF = tab_floats(w, h);
		int lineRes = 0;
		for (int i = 0; i < nViews; i++) {
			fill_view(scenes[i], w, h, under, fogDist, T->h_views[i]);
			lineRes = T->h_views[i].lineRes;
		}
		// ---- chunking: strips of one chunk must fit the budget; chunks are made equal (the two streams overlap
		// best with equal partners), and a large batch is split in two even when it would fit one chunk ----
		size_t budgetLines = std::max<size_t>(strip_budget_bytes() / ((size_t)lineRes * 8), 65536);
		{
			size_t total = 0;
			for (int i = 0; i < nViews; i++)
				total += (size_t)T->h_views[i].numLines;
			size_t want = (total + budgetLines - 1) / budgetLines;
			if (want < 2 && nViews >= 512 && dual_stream())
				want = 2;
			// frames going back to the host: the PCIe copy (4-8 B/pixel) is the long pole, so start it early --
			// more, smaller chunks cost kernel time that hides under the copy of the previous chunk
			if ((host_fb || host_depth) && nViews >= 1024 && want < 8)
				want = 8;
			budgetLines = std::max<size_t>((total + want - 1) / want + 65536 / 32, 65536);
		}
		std::vector<int> chunkStart;
		size_t maxLines = 0, maxViews = 0;
		{
			size_t lines = 0;
			int start = 0;
			chunkStart.push_back(0);
			for (int i = 0; i < nViews; i++) {
				ViewParams &v = T->h_views[i];
				if (lines + (size_t)v.numLines > budgetLines && i > start) {
					chunkStart.push_back(i);
					start = i;
					lines = 0;
				}
				v.lineBase = (uint32_t)lines;
				v.tabBase = (uint32_t)(i - start) * tabF;
				lines += (size_t)v.numLines;
				maxLines = std::max(maxLines, lines);
				maxViews = std::max(maxViews, (size_t)(i - start + 1));
			}
			chunkStart.push_back(nViews);
		}
		const bool hostPrep = nViews <= 8;
		const size_t nChunks = chunkStart.size() - 1;
		// one chunk (or the host-prep path with its shared pinned staging): everything on the context stream
		// (per-kernel event timing wants each kernel alone on the GPU: with timing on, chunks run back to back)
		const bool dual = dual_stream() && nChunks > 1 && !hostPrep && !c->timing;
		for (int k = 0; k < (dual ? 2 : 1); k++) {
			ChunkBufs &B = T->set[k];
			if (grow(B.d_lineDirs, B.lineCap, maxLines) || grow(B.d_lineParams, B.lineParamCap, maxLines) ||
			    grow(B.d_strips, B.stripCap, maxLines * (size_t)lineRes) ||
			    grow(B.d_lineMask, B.maskCap, maxLines * (size_t)(lineRes >> 5)) || grow(B.d_tabs, B.tabCap, maxViews * tabF))
				return -2;
		}
		OSB_CUDA(cudaMemcpyAsync(T->d_views, T->h_views, sizeof(ViewParams) * nViews, cudaMemcpyHostToDevice,
		                         c->stream));
		if (hostPrep) {
			if (grow(T->h_lineDirs, T->hLineCap, maxLines, true))
				return -2;
			if (grow(T->h_tabs, T->hTabCap, maxViews * tabF, true))
				return -2;
		}
		if (dual) {
			OSB_CUDA(cudaEventRecord(T->fork, c->stream));
			for (ChunkBufs &B : T->set)
				OSB_CUDA(cudaStreamWaitEvent(B.stream, T->fork, 0));
		}
		osbn::SseTables sse{c->d_rcp, c->d_rsqrt, c->rcpBits, c->rsqrtBits};
		const int num = c->opts.numerics;
		for (size_t ci = 0; ci < nChunks; ci++) {
			ChunkBufs &B = T->set[dual ? (ci & 1) : 0];
			const cudaStream_t st = dual ? B.stream : c->stream;
			const int v0 = chunkStart[ci], nv = chunkStart[ci + 1] - v0;
			const ViewParams &last = T->h_views[v0 + nv - 1];
			const uint32_t totalLines = last.lineBase + (uint32_t)last.numLines;
			const ViewParams *dV = T->d_views + v0;
			if (hostPrep) {
				for (int i = 0; i < nv; i++) {
					const ViewParams &V = T->h_views[v0 + i];
					float *vx = T->h_tabs + V.tabBase, *vy = vx + (w >> 2), *v1 = vy + (h >> 2);
					prep_line_dirs(V, T->h_lineDirs + V.lineBase);
					prep_fog_vx(V, g, vx);
					for (int t = 0; t < g.swThreads; t++) {
						prep_fog_vy(V, g, t, vy);
						prep_v1(V, g, t, v1);
					}
				}
				OSB_CUDA(cudaMemcpyAsync(B.d_lineDirs, T->h_lineDirs, sizeof(float2) * totalLines,
				                         cudaMemcpyHostToDevice, st));
				OSB_CUDA(cudaMemcpyAsync(B.d_tabs, T->h_tabs, sizeof(float) * tabF * nv, cudaMemcpyHostToDevice, st));
			} else {
				KLaunch kl(c, K_VIEW_PREP, st);
				k_view_prep<<<(nv + 3) / 4, 128, 0, st>>>(dV, nv, g, B.d_lineDirs, B.d_tabs);
			}
			// small launches use fewer lanes per warp (a line is one long dependent chain: what counts is how many
			// warps run side by side and how little each warp diverges)
			int lpw = 32;
			while (lpw > line_min_lanes() && (size_t)totalLines / lpw < (size_t)c->smCount * 4 * 2) // < 2 warps per scheduler
				lpw >>= 1;
			const unsigned linesPerCta = kLineThreads / 32 * lpw;
			const unsigned lineBlocks = (totalLines + linesPerCta - 1) / linesPerCta;
			size_t smem = (size_t)(lineRes >> 5) * kLineThreads * sizeof(uint32_t);
			if (OSB_RCP_SMEM && num != OSB200_NUM_IEEE && c->rcpBits <= kRcpSmemBits)
				smem += sizeof(uint32_t) << c->rcpBits;
#define LAUNCH_LINES(SSE, SH)                                                                      \
	k_build_lines<SSE, SH><<<lineBlocks, kLineThreads, smem, st>>>(                                 \
	  dV, nv, B.d_lineDirs, c->d_solid, c->d_color, sse, B.d_strips, B.d_lineParams, B.d_lineMask, totalLines, lineRes, lpw)
			if (!g.noMap) {
				KLaunch kl(c, K_BUILD_LINES, st);
				if (num == OSB200_NUM_IEEE)
					LAUNCH_LINES(false, false);
				else if (num == OSB200_NUM_SSE)
					LAUNCH_LINES(true, false);
				else
					LAUNCH_LINES(true, true);
			}
#undef LAUNCH_LINES
			const int segs = ((h >> 3) + kSegRows - 1) / kSegRows;
			const size_t tasks = (size_t)nv * (((w >> 3) + 3) >> 2) * segs;
			const unsigned gBlocks = (unsigned)((tasks + kGatherWarps - 1) / kGatherWarps);
			uint32_t *fbC = d_fb + (size_t)v0 * w * h;
			float *dbC = d_depth + (size_t)v0 * w * h;
#define LAUNCH_GATHER(SSE, FOG)                                                                    \
	k_gather_fog<SSE, FOG><<<gBlocks, kGatherWarps * 32, 0, st>>>(                                  \
	  dV, nv, g, B.d_tabs, c->d_solid, T->d_atan, sse, B.d_strips, B.d_lineParams, B.d_lineMask, lineRes, fbC, dbC)
			{
				KLaunch kl(c, K_GATHER_FOG, st);
				if (num == OSB200_NUM_IEEE)
					LAUNCH_GATHER(false, false);
				else if (num == OSB200_NUM_SSE)
					LAUNCH_GATHER(true, false);
				else
					LAUNCH_GATHER(true, true);
			}
#undef LAUNCH_GATHER
			OSB_CUDA(cudaEventRecord(B.done, st));
			if (host_fb || host_depth) {
				const size_t off = (size_t)v0 * w * h, cnt = (size_t)nv * w * h;
				OSB_CUDA(cudaStreamWaitEvent(T->copyStream, B.done, 0));
				if (host_fb)
					OSB_CUDA(cudaMemcpyAsync(host_fb + off, d_fb + off, cnt * 4, cudaMemcpyDeviceToHost, T->copyStream));
				if (host_depth)
					OSB_CUDA(cudaMemcpyAsync(host_depth + off, d_depth + off, cnt * 4, cudaMemcpyDeviceToHost, T->copyStream));
			}
			if (hostPrep && ci + 1 < nChunks) // staging reused by the next chunk
				OSB_CUDA(cudaStreamSynchronize(st));
		}
		if (dual) { // join: the context stream continues after both working streams
			for (ChunkBufs &B : T->set)
				OSB_CUDA(cudaStreamWaitEvent(c->stream, B.done, 0));
		}
		OSB_CUDA(cudaGetLastError());
		OSB_CUDA(cudaEventRecord(T->stageFree, c->stream));
		return 0;
	}
	// Context-owned frame + depth of the host-pointer entry points.  The depth buffer persists between frames like
	// SWRenderer::depthBuffer (never cleared, SWRenderer.cpp:83); a fresh one starts at zero like vector::resize.
	int terrain_out_buffers(osb200_ctx *c, size_t n, uint32_t **fb, float **depth) {
		if (terrain_init(c))
			return -2;
		TerrainState *T = c->terrain;
		if (n != T->outCap) {
			cudaFree(T->d_fb);
			cudaFree(T->d_depth);
			T->d_fb = nullptr;
			T->d_depth = nullptr;
			T->outCap = 0;
			OSB_CUDA(cudaMalloc(&T->d_fb, n * 4));
			OSB_CUDA(cudaMalloc(&T->d_depth, n * 4));
			OSB_CUDA(cudaMemsetAsync(T->d_depth, 0, n * 4, c->stream));
			T->outCap = n;
		}
		// the previous call's copies must not race with this call's kernels overwriting the same device buffers
		OSB_CUDA(cudaStreamSynchronize(T->copyStream));
		*fb = T->d_fb;
		*depth = T->d_depth;
		return 0;
	}

	// ApplyFog over the single view last set up by render_impl(.., deferFog = true) on the context stream
	// The tail of SWRenderer::EndScene for a frame that drew models and / or dynamic lights, in one pass over the frame:
	//   winner of the splat keys (SWModelRenderer's depth-tested writes, models.cu)      keys == null: nothing was splatted
	//   ApplyDynamicLight (SWRenderer.cpp:283-325) for every kept light in submission order, float recurrences from
	//   the tables of k_light_prep                                                       nLights == 0: no lights
	//   ApplyFog (SWRenderer.cpp:341-516), same per-4x4-block scale and blend as k_gather_fog
	template <bool kSse, bool kSse2Fog>
	__global__ void __launch_bounds__(256)
	  k_scene_finish(const ViewParams *__restrict__ views, FrameGeom g, const float *__restrict__ tabs, osbn::SseTables sse,
	                 uint32_t *__restrict__ fb, float *__restrict__ depthBuf, const unsigned long long *__restrict__ keys,
	                 const uint32_t *__restrict__ splatColor, const LightDev *__restrict__ lights, int nLights,
	                 const float *__restrict__ lightTabs) {
		const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
		if (x >= g.fw || y >= g.fh)
			return;
		const size_t o = (size_t)y * g.fw + x;
		uint32_t color;
		float z;
		bool splat = false;
		if (keys) {
			const unsigned long long k = keys[o];
			const uint32_t seq = (uint32_t)k;
			if (seq) {
				color = splatColor[seq];
				z = key_depth((uint32_t)(k >> 32));
				depthBuf[o] = z;
				splat = true;
			}
		}
		if (!splat) {
			color = fb[o];
			z = depthBuf[o];
		}
		for (int li = 0; li < nLights; li++) {
			const LightDev &L = lights[li];
			if (x < L.minX || x >= L.maxX || y < L.minY || y >= L.maxY)
				continue;
			const float *vxT = lightTabs + L.tabBase, *vyT = vxT + g.fw;
			const float px = __fsub_rn(__fmul_rn(vxT[x], z), L.center[0]);
			const float py = __fsub_rn(__fmul_rn(vyT[y], z), L.center[1]);
			const float pz = __fsub_rn(z, L.center[2]);
			float dist = __fadd_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)), __fmul_rn(pz, pz));
			dist = __fmul_rn(dist, L.invRadius2);
			if (dist < 1.f) {
				float strength = __fsub_rn(1.f, dist);
				strength = __fmul_rn(strength, strength);
				strength = __fmul_rn(strength, 256.f);
				const int factor = osbn::x86_f2i(strength);
				const uint32_t sr = (color >> 16) & 0xff, sg = (color >> 8) & 0xff, sb = color & 0xff;
				const int ar = (int)((uint32_t)(L.r * factor) * sr), ag = (int)((uint32_t)(L.g * factor) * sg),
				          ab = (int)((uint32_t)(L.b * factor) * sb);
				const uint32_t dr = min((uint32_t)(ar >> 16) + sr, 255u), dg = min((uint32_t)(ag >> 16) + sg, 255u),
				               db = min((uint32_t)(ab >> 16) + sb, 255u);
				color = db | (dg << 8) | (dr << 16);
			}
		}
		const ViewParams &V = views[0];
		osbn::Num<kSse> num{sse};
		const float *vxTab = tabs + V.tabBase, *vyTab = vxTab + (g.fw >> 2);
		const float vx = vxTab[x >> 2], vy = vyTab[y >> 2];
		float ds = __fadd_rn(__fadd_rn(1.f, __fmul_rn(vx, vx)), __fmul_rn(vy, vy));
		ds = __fmul_rn(ds, __fmul_rn(num.rsqrt(ds), V.fogScale));
		const uint32_t fogBR = g.fogB | (g.fogR << 16), fogG8 = g.fogG << 8;
		fb[o] = fog_blend<kSse2Fog>(color, __fmul_rn(z, ds), fogBR, fogG8);
	}

	int terrain_finish_scene(osb200_ctx *c, int w, int h, const float fogColor[3], uint32_t *d_fb, float *d_depth,
	                         const unsigned long long *d_keys, const uint32_t *d_splatColor, const LightDev *d_lights,
	                         int nLights, const float *d_lightTabs) {
		TerrainState *T = c->terrain;
		FrameGeom g;
		std::memset(&g, 0, sizeof(g));
		g.fw = w;
		g.fh = h;
		g.under = c->opts.undersampling;
		g.swThreads = c->opts.sw_threads;
		g.sse2Fog = g.sse2Shade = c->opts.numerics == OSB200_NUM_SSE2;
		auto toFixed8 = [](float v) {
			int i = static_cast<int>(v * 255.f + .5f);
			return (uint32_t)std::max(std::min(i, 255), 0);
		};
		g.fogR = toFixed8(fogColor[0]);
		g.fogG = toFixed8(fogColor[1]);
		g.fogB = toFixed8(fogColor[2]);
		osbn::SseTables sse{c->d_rcp, c->d_rsqrt, c->rcpBits, c->rsqrtBits};
		const dim3 grid((w + 31) / 32, (h + 7) / 8, 1);
		KLaunch kl(c, K_SCENE_FINISH);
#define LAUNCH_FINISH(SSE, FOG)                                                                                        \
	k_scene_finish<SSE, FOG><<<grid, 256, 0, c->stream>>>(T->d_views, g, T->set[0].d_tabs, sse, d_fb, d_depth, d_keys,   \
	                                                      d_splatColor, d_lights, nLights, d_lightTabs)
		if (c->opts.numerics == OSB200_NUM_IEEE)
			LAUNCH_FINISH(false, false);
		else if (c->opts.numerics == OSB200_NUM_SSE)
			LAUNCH_FINISH(true, false);
		else
			LAUNCH_FINISH(true, true);
#undef LAUNCH_FINISH
		OSB_CUDA(cudaGetLastError());
		return 0;
	}

} // namespace osb

extern "C" {
#ifdef OSB_STATS
__attribute__((visibility("default"))) int osb200_debug_stats(unsigned long long *out, int reset) {
	cudaDeviceSynchronize();
	cudaMemcpyFromSymbol(out, osb::g_stats, sizeof(unsigned long long) * 16);
	if (reset) {
		unsigned long long z[16] = {0};
		cudaMemcpyToSymbol(osb::g_stats, z, sizeof(z));
	}
	return 0;
}
#endif

int osb200_render_device(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h,
                         const float fogColor[3], float fogDist, uint32_t *d_bgrx, float *d_depth) {
	return osb::render_impl(c, scenes, nViews, w, h, fogColor, fogDist, d_bgrx, d_depth, nullptr, nullptr, false);
}

int osb200_render(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
                  float fogDist, uint32_t *host_bgrx, float *host_depth) {
	if (!c) {
		osb::set_error("osb200_render: no context");
		return -1;
	}
	if (nViews <= 0)
		return nViews == 0 ? 0 : -1;
	if (w <= 0 || h <= 0 || (w & 7) || (h & 7)) {
		osb::set_error("Framebuffer size is not multiple of 8.");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::terrain_init(c))
		return -2;
	osb::TerrainState *T = c->terrain;
	uint32_t *ofb;
	float *odp;
	if (int rc0 = osb::terrain_out_buffers(c, (size_t)nViews * w * h, &ofb, &odp))
		return rc0;
	int rc = osb::render_impl(c, scenes, nViews, w, h, fogColor, fogDist, T->d_fb, T->d_depth, host_bgrx, host_depth, false);
	if (rc)
		return rc;
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	OSB_CUDA(cudaStreamSynchronize(T->copyStream));
	return 0;
}

int osb200_read_depth(osb200_ctx *c, float *out, size_t n) {
	if (!c || !out) {
		osb::set_error("osb200_read_depth: null argument");
		return -1;
	}
	if (!c->terrain || n > c->terrain->outCap) {
		osb::set_error("osb200_read_depth: no frame of that size has been rendered by osb200_render");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	OSB_CUDA(cudaMemcpyAsync(out, c->terrain->d_depth, n * 4, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

} // extern "C"
