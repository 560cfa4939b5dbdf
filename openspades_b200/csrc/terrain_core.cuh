// Per-lane core of the terrain pass, shared by the CUDA kernels (terrain.cu) and by the CPU lock-step emulation of
// those kernels (tests/native/terrain_emul.cpp).  Everything here is plain per-thread code: the kernels add the
// warp-level scheduling (work cursors, lane refill, __syncwarp) around it, the emulation runs the lanes one after the
// other.  Reference (paths relative to /root/reference/Sources):
//   SWMapRenderer::RenderInner   Draw/SWMapRenderer.cpp:1045-1174  per-view set-up (fill_view in terrain.cu + prep_*)
//   SWMapRenderer::BuildLine     Draw/SWMapRenderer.cpp:219-671    line_setup (:232-305) + LineWalker (:309-671)
//   SWMapRenderer::RenderFinal   Draw/SWMapRenderer.cpp:744-1043   GatherLane
//   SWRenderer::ApplyFog         Draw/SWRenderer.cpp:341-516       fog_blend, fused into GatherLane
// All float arithmetic is single IEEE operations in the reference's evaluation order (osbc:: wrappers, hd.cuh).
#pragma once
#include "../../include/osb200.h"
#include "hd.cuh"
#include "libm_f32.cuh"
#include "numerics.cuh"

#include <cmath>
#include <cstdlib>
#include <cstring>

#ifndef OSB_STAT
#define OSB_STAT(what, n) // emulation-only counters (tests/native/terrain_emul.cpp)
#endif
#define OSB_PRAGMA_(x) _Pragma(#x)
#define OSB_PRAGMA(x) OSB_PRAGMA_(x)
#ifndef OSB_GATHER_HALF_UNROLL
#define OSB_GATHER_HALF_UNROLL 1 // 1 = the two 4-row halves of a block row share one copy of the pixel code; 2 = unrolled
#endif
#ifndef OSB_STRIP_HULL
#define OSB_STRIP_HULL 1 // 1: a finished line publishes the hull [first, last] of its written strip pixels in its LineRec and
                         //    stores LinePixel::Clear into the (rare) holes inside it: the gather needs no mask word;
                         // 0: the line publishes its written-pixel bitmask, the gather tests the bit before loading a pixel
#endif
#ifndef OSB_GATHER_BATCH
#define OSB_GATHER_BATCH 4 // pixel rows whose loads are in flight together (4 or 8)
#endif
#ifndef OSB_GATHER_ROW_UNROLL
#define OSB_GATHER_ROW_UNROLL 1 // same for the two block rows of a pass
#endif

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
		uint32_t lineBase; // first line slot of this view inside the chunk
		uint32_t tabBase;  // first float of this view's recurrence tables
		float fogScale;    // 255 / fogDistance
		float dvx, dvy;    // ApplyFog steps
		uint32_t pad;
	};

	// Recurrence tables of one view, laid out at tabs + tabBase:
	//   vx[fw/4]            ApplyFog per-4-pixel-column vx          (SWRenderer.cpp:372,409)
	//   vy[fh/4]            ApplyFog per-4-row vy, per emulated thread band (:365-369,412)
	//   v1[fw/8][3]         RenderFinal per-8-pixel-column start vector, per emulated thread (:757,789,1040)
	OSB_HDI uint32_t tab_floats(int fw, int fh) { return fw / 4 + fh / 4 + (fw / 8) * 3; }

	struct FrameGeom {
		int fw, fh, under, swThreads, sse2Fog, sse2Shade;
		int noFog; // fog deferred to k_scene_finish (models / dynamic lights are drawn between terrain and fog, SWRenderer.cpp:727-744)
		int noMap; // no map set: SWRenderer::EndScene still clears and fogs over the stale depth (SWRenderer.cpp:713-744)
		uint32_t fogB, fogG, fogR;
	};

	// The solid bitmap is stored in 4x4-column tiles of 128 B (see ctx.h: tiled_index; repeated here so that the
	// core does not depend on the context header).
	OSB_HDI uint32_t tiled_xy(uint32_t x, uint32_t y) { return ((x >> 2) * 128u + (y >> 2)) * 16u + (x & 3u) * 4u + (y & 3u); }

	// One sequential float recurrence per role; the same code runs on the host (few views) or on one
	// lane per role on the device (many views).
	OSB_HDI void prep_line_dirs(const ViewParams &v, float2 *dirs) {
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
	OSB_HDI void prep_fog_vx(const ViewParams &v, const FrameGeom &g, float *vx) {
		float x = v.fovXt;
		for (int i = 0; i < g.fw / 4; i++) {
			vx[i] = x;
			x = osbm::addf(x, v.dvx);
		}
	}
	OSB_HDI void prep_fog_vy(const ViewParams &v, const FrameGeom &g, int t, float *vy) {
		// band of emulated thread t, SWRenderer.cpp:359-369
		int startY = (int)((unsigned)g.fh * (unsigned)t / (unsigned)g.swThreads) & ~3;
		int endY = (int)((unsigned)g.fh * (unsigned)(t + 1) / (unsigned)g.swThreads) & ~3;
		float y = osbm::addf(v.fovYt, osbm::mulf(v.dvy, (float)(startY >> 2)));
		for (int r = startY; r < endY; r += 4) {
			vy[r >> 2] = y;
			y = osbm::addf(y, v.dvy);
		}
	}
	OSB_HDI void prep_v1(const ViewParams &v, const FrameGeom &g, int t, float *v1tab) {
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

	// Host half of RenderInner (SWMapRenderer.cpp:1057-1133) + BuildFrustrum (SWRenderer.cpp:207-227) +
	// ApplyFog's scalars (SWRenderer.cpp:345-357) for one view, with the host libm (host-only: no __device__).
	inline void fill_view(const osb200_scene &s, int fw, int fh, int under, float fogDist, ViewParams &v) {
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

	// ------------------------------------------------------------------------------------------------------------
	// Per-line records.  k_line_setup writes one LineRec per yaw line of the chunk (culled or not) plus the float
	// pair the exact per-pixel path of the gather needs; lines that survive the pitch cull are appended to the
	// chunk's active list, and their position in that list is their strip slot.
	// ------------------------------------------------------------------------------------------------------------
	struct alignas(16) LineRec {
		int pitchTanMinI; // static_cast<int>(minTan * 65536.f)        (SWMapRenderer.cpp:305)
		int pitchScaleI;  // static_cast<int>(pitchScale * 65536.f)    (:306)
		int slot;         // strip slot (index in the active list), -1 = culled line: its strip is all LinePixel::Clear;
		                  // hull scheme: slot * lineRes, the index of the line's first strip pixel (make_line_rec)
		int view;         // view index inside the chunk until the walker has finished the line; then (OSB_STRIP_HULL)
		                  // first | count << 16: the hull [first, first + count) of the strip pixels the line wrote
	};
	constexpr int kEmptyHull = 0; // count = 0
	static_assert(sizeof(LineRec) == 16, "LineRec is loaded as one int4");

	struct LineSetupOut {
		float minTan, pitchScale;
		bool culled;
	};

	// Pitch range of one yaw line from the four frustum side planes, SpecialTan of its ends, pitch scale
	// (SWMapRenderer.cpp:232-305).  hd = the line's horizonDir.
	template <bool kSse> OSB_HDI LineSetupOut line_setup(const ViewParams &V, float2 hd, int lineRes, const osbn::Num<kSse> &num) {
		using namespace osbc;
		const float pi = 3.14159274101257324f; // static const float pi = M_PI
		float minPitch = -pi * 0.4999f, maxPitch = pi * 0.4999f;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
		for (int k = 0; k < 4; k++) {
			const float px = V.frN[k][0], py = V.frN[k][1], pz = V.frN[k][2];
			if (px == 0.f && py == 0.f) {
				if (pz > 0.f)
					minPitch = minPitch < 0.f ? 0.f : minPitch;
				else
					maxPitch = 0.f < maxPitch ? 0.f : maxPitch;
			} else if (pz == 0.f) {
				// Vector3::Dot(plane, horz) with horz.z = 0: x*x' + y*y' + z*0
				const float d = fadd(fadd(fmul(px, hd.x), fmul(py, hd.y)), fmul(pz, 0.f));
				if (d < 0.f) {
					minPitch = 2.f;
					maxPitch = -2.f;
				}
			} else {
				// prj = (px,py,0).Normalize()
				float sc = fsqrt(fadd(fadd(fmul(px, px), fmul(py, py)), 0.f));
				if (sc != 0.f)
					sc = fdiv(1.f, sc);
				const float qx = fmul(px, sc), qy = fmul(py, sc), qz = fmul(0.f, sc);
				const float zv = fabsf(pz);
				const float cs = fadd(fadd(fmul(qx, hd.x), fmul(qy, hd.y)), fmul(qz, 0.f));
				const float cs2 = fmul(cs, cs);
				float ang = fdiv(fmul(fmul(zv, zv), fsub(1.f, cs2)), cs2);
				ang = fmul(-cs, num.sqrt(fadd(1.f, ang)));
				ang = fdiv(zv, ang);
				if (isnan_(ang) || isinf_(ang) || ang == 0.f)
					continue;
				ang = fdiv(num.sqrt(fsub(1.f, fmul(ang, ang))), ang);
				ang = osbm::atanf_glibc(ang);
				if (isnan_(ang) || isinf_(ang))
					continue;
				if (pz > 0.f) {
					const float c = fsub(ang, 0.01f);
					minPitch = minPitch < c ? c : minPitch;
				} else {
					const float c = fadd(-ang, 0.01f);
					maxPitch = c < maxPitch ? c : maxPitch;
				}
			}
		}
		if (minPitch >= maxPitch) { // culled (:387-389): the strip stays cleared and nothing reads the line's pitch mapping --
			LineSetupOut c;         // its record carries the empty hull, so the two tanf below would be computed for nobody
			c.minTan = 0.f;
			c.pitchScale = 0.f;
			c.culled = true;
			return c;
		}
		// SpecialTan (:43-57)
		auto specialTan = [&](float v) -> float {
			if (v <= -pi * 0.5f)
				return -2.f;
			else if (v < -pi * 0.25f)
				return fsub(-2.f, fdiv(1.f, osbm::tanf_glibc(v)));
			else if (v < pi * 0.25f)
				return osbm::tanf_glibc(v);
			else if (v < pi * 0.5f)
				return fsub(2.f, fdiv(1.f, osbm::tanf_glibc(v)));
			return 2.f;
		};
		const float minTan = specialTan(minPitch);
		float maxTan = specialTan(maxPitch);
		{
			const float minDiff = fdiv((float)lineRes, 10000.f);
			if (maxTan < fadd(minTan, minDiff))
				maxTan = fadd(minTan, minDiff);
		}
		LineSetupOut o;
		o.minTan = minTan;
		o.pitchScale = fdiv((float)lineRes, fsub(maxTan, minTan));
		o.culled = minPitch >= maxPitch; // the strip stays cleared (:387-389)
		return o;
	}
	// lineRes > 0 (hull scheme): the record carries the index of the line's first strip pixel, slot * lineRes, instead of
	// the slot -- the gather adds the pitch index and is done (one multiply and one constant load less per pixel)
	OSB_HDI LineRec make_line_rec(const LineSetupOut &o, int slot, int view, int lineRes = 0) {
		LineRec r;
		r.pitchTanMinI = osbn::x86_f2i(osbc::fmul(o.minTan, 65536.f));
		r.pitchScaleI = osbn::x86_f2i(osbc::fmul(o.pitchScale, 65536.f));
		r.slot = (OSB_STRIP_HULL && lineRes > 0 && slot >= 0) ? (int)((uint32_t)slot * (uint32_t)lineRes) : slot;
		r.view = view;
		return r;
	}

	// How many DDA steps a line will take, estimated at set-up -- ORDER of the walker's work only, no effect on any
	// result.  A line ends at the fog distance (128) or where the every-4th-step pitch cull fires (:661-667): once the
	// map's top layer (z = 0) projects at or above the last strip pixel, or its bottom layer (z = 63) at or below the
	// first.  Both projections are monotonic in the distance, so the distance of the cull has a closed form (the inverse
	// of ToSpecialTan, :59-66); steps = distance x cell crossings per unit distance (|dx| + |dy| of the unit direction).
	constexpr int kTaskBuckets = 16;      // walker tasks are queued by estimated length, longest first
	constexpr float kStepsPerBucket = 12.f; // 128 * sqrt(2) = 181 steps at most
	OSB_HDI float predict_steps(float minTan, float pitchScale, int lineRes, float cz, float2 hd) {
		float dEnd = 128.f;
		if (pitchScale > 0.f) {
			const float inv = 1.f / pitchScale;
			const float tA = (float)(lineRes - 1) * inv + minTan, tB = inv + minTan;
			if (fabsf(tA) < 2.f && cz > 0.f) { // z = 0 lies above the eye: v = -cz / d rises towards 0
				const float vA = fabsf(tA) <= 1.f ? tA : 1.f / ((tA < 0.f ? -2.f : 2.f) - tA);
				if (vA < 0.f)
					dEnd = fminf(dEnd, cz / -vA);
			}
			if (fabsf(tB) < 2.f && cz < 63.f) { // z = 63 lies below the eye: v = (63 - cz) / d falls towards 0
				const float vB = fabsf(tB) <= 1.f ? tB : 1.f / ((tB < 0.f ? -2.f : 2.f) - tB);
				if (vB > 0.f)
					dEnd = fminf(dEnd, (63.f - cz) / vB);
			}
		}
		return dEnd * (fabsf(hd.x) + fabsf(hd.y));
	}
	OSB_HDI int task_bucket(float steps) {
		const int b = (int)(steps * (1.f / kStepsPerBucket));
		return b < 0 ? 0 : (b >= kTaskBuckets ? kTaskBuckets - 1 : b);
	}

	// BuildLinePixel colour (SWMapRenderer.cpp:532-576). col = 0xHHBBGGRR from the map; mode 0 = NegY walls and
	// tops (NegZ), 1 = PosX/PosY/NegX walls, 2 = undersides (PosZ).  All four 8-bit lanes are shaded in one
	// register: (x>>1)+(x>>2) <= 190 never carries into the neighbouring lane.
	template <bool kSse2Shade> OSB_HDI uint32_t shade(uint32_t col, int mode) {
		if (kSse2Shade) {
			// 16-bit lanes B,G,R,H: PosZ >>1; PosX/PosY/NegX (>>1)+(>>2); health<100 >>1; pack; filled bit
			const bool weak = (col >> 24) < 100u;
			const uint32_t half = (col >> 1) & 0x7f7f7f7fu, quarter = (col >> 2) & 0x3f3f3f3fu;
			uint32_t v = mode == 2 ? half : (mode == 1 ? half + quarter : col);
			v = weak ? (v >> 1) & 0x7f7f7f7fu : v;
			return osbc::byte_perm(v, 0u, 0x3012) | (1u << 24); // H B G R -> H R G B
		} else {
			uint32_t c = osbc::byte_perm(col, 0u, 0x4012); // 0 R G B
			c = mode == 2 ? (c & 0xfcfcfcu) >> 2 : (mode == 1 ? (c & 0xfefefeu) >> 1 : c);
			return c | (1u << 24);
		}
	}

	// What a walker lane reads and writes; one per launch (kernel parameters) / per emulated chunk.
	struct WalkIO {
		const ViewParams *views;
		const float2 *lineDirs;   // horizonDir per line of the chunk
		const float2 *lineTan;    // {pitchTanMin, pitchScale} per line of the chunk
		LineRec *lineRecs;        // read at init (view); the hull is written back when the line ends
		const uint64_t *solid;    // [tiled_xy(x, y)]
		const uint32_t *color;    // [tiled_xy(x, y) * 64 + z]
		uint2 *strips;            // [slot * lineRes + i]  LinePixel {combined, depth bits}
		uint32_t *lineMask;       // [slot * (lineRes / 32) + w] strip pixels the line wrote
		int lineRes;
	};

	// ------------------------------------------------------------------------------------------------------------
	// One yaw line (SWMapRenderer::BuildLine from the DDA set-up on, SWMapRenderer.cpp:309-671), as a resumable
	// per-lane state machine: init() once, then step() until it returns false.
	//   * the reference's six per-column RLE z-lists are never stored: tops = s & ~(s<<1), bottoms = s & ~(s>>1),
	//     walls = sNew & ~sOld of the 64-bit column masks of the cell left / entered, walked lowest bit first in
	//     the reference's list order;
	//   * "first writer wins" is a bitmask of the strip (`filled`, one word per 32 strip pixels plus four words of
	//     eye masks = kWalkExtraWords behind them, in shared memory on the device), so strip pixels are written once and colours are fetched only for spans that land; the strip
	//     is never cleared: the mask is published per line and the gather substitutes LinePixel::Clear;
	//   * the DDA keeps only what later steps depend on: the cell and the in-cell fraction MIRRORED into the
	//     direction of travel (g = f for a positive direction, 511 - f for a negative one), which turns the
	//     reference's four sign cases (:432-522) into one, with identical integers;
	//   * the step after the current one is computed ahead and its column mask requested before the current
	//     step's spans are processed: the dependent 8-byte load is off the critical path.
	// ------------------------------------------------------------------------------------------------------------
	constexpr int kWalkExtraWords = 4; // per-lane words behind the fill mask: belowEye lo/hi, aboveEye lo/hi
	template <bool kSse, bool kSse2Shade> struct LineWalker {
		// DDA
		uint32_t invDirXI, invDirYI, dirXI, dirYI;
		int irx, iry;       // current cell (unwrapped)
		uint32_t gx, gy;    // mirrored in-cell fraction, 0..511
		uint32_t dirBits;   // bit0 posX, bit1 posY
		float distance, invDist;
		// look-ahead step
		int nIrx, nIry;
		uint32_t nGx, nGy, nT;
		float nDistance;
		int nWallMode;
		uint64_t sOld, sNext;
		uint32_t tOld;
		// projection
		float cz, transScale, transOffset, zscale, heightScale, depthBias;
		int icz;
		uint32_t stripBase, maskBase; // slot * lineRes; slot * words (mask scheme) or the line's index (hull scheme)
		uint32_t cnt2;
		int count;

		static constexpr float kFogDist = 128.f;

		// transform lambda, SWMapRenderer.cpp:395-401 (ToSpecialTan :59-66 inlined)
		OSB_HDM int transform(const osbn::Num<kSse> &num, float iDist, int z, float vmax) const {
			using namespace osbc;
			const float zv = fsub(i2f(z), cz);
			const float v = fmul(iDist, zv);
			float t = v;
			if (fabsf(v) > 1.f) // |v| <= 64e20: a normal number, the table path needs no range test
				t = num.specialTanTail(v);
			float p = fadd(fmul(t, transScale), transOffset);
			// p < 0 ? 0 : p;  vmax < p ? vmax : p;  truncation.  p is never NaN here (finite * finite), and a NaN
			// would end as 0 either way (cvttss2si of the clamp result vs fmaxf dropping the NaN).
			// vmax = lineRes + 0.5: trunc(min(max(p, 0), vmax)) == min(trunc_sat_unsigned(p), lineRes) -- the saturating
			// unsigned conversion is the lower clamp, and an integer min the upper one
			return (int)uminu(f2u_rz_sat(p), (uint32_t)vmax);
		}

		// one DDA step from (irx0, iry0, gx0, gy0, dist0) into the n* members (:426-522)
		OSB_HDM void advance(int irx0, int iry0, uint32_t gx0, uint32_t gy0, float dist0) {
			using namespace osbc;
			const uint32_t posX = dirBits & 1u, posY = (dirBits >> 1) & 1u;
			const uint32_t tx = (511u + posX - gx0) * invDirXI;
			const uint32_t ty = (511u + posY - gy0) * invDirYI;
			const bool xs = tx < ty;
			const uint32_t tt = xs ? tx : ty;
			const uint32_t d = ((xs ? dirYI : dirXI) * tt) >> 17;
			nIrx = irx0 + (xs ? (posX ? 1 : -1) : 0);
			nIry = iry0 + (xs ? 0 : (posY ? 1 : -1));
			nGx = xs ? 0u : ((gx0 + d) & 511u);
			nGy = xs ? ((gy0 + d) & 511u) : 0u;
			nDistance = fadd(dist0, fmul(u2flt(tt), 1.f / 512.f / 256.f));
			// wall face of an x crossing is NegX / PosX (both 3/4-shaded, mode 1); of a y crossing NegY (mode 0) / PosY (1)
			nWallMode = xs ? 1 : (posY ? 0 : 1);
			nT = tiled_xy((uint32_t)nIrx & 511u, (uint32_t)nIry & 511u);
		}

		// z > icz / z < icz as bit masks (the eye's voxel layer splits a column into floors seen from above and ceilings
		// seen from below)
		OSB_HDM void eye_masks(uint32_t m[4]) const {
			const uint64_t belowEye = icz >= 63 ? 0ull : (icz < 0 ? ~0ull : (~0ull << (icz + 1)));
			const uint64_t aboveEye = icz <= 0 ? 0ull : (icz >= 64 ? ~0ull : ((1ull << icz) - 1ull));
			m[0] = (uint32_t)belowEye;
			m[1] = (uint32_t)(belowEye >> 32);
			m[2] = (uint32_t)aboveEye;
			m[3] = (uint32_t)(aboveEye >> 32);
		}

		// everything of the set-up that does not touch the fill mask or the map (shared with the cooperative walker)
		OSB_HDM void init_consts(const WalkIO &io, uint32_t gl, uint32_t slot) {
			using namespace osbc;
			const LineRec rec = io.lineRecs[gl];
			const ViewParams &V = io.views[rec.view];
			const float2 hd = io.lineDirs[gl];
			const float2 tn = io.lineTan[gl];
			const int lineRes = io.lineRes;
			stripBase = slot * (uint32_t)lineRes;
#if OSB_STRIP_HULL
			maskBase = gl;
#else
			maskBase = slot * (uint32_t)(lineRes >> 5);
#endif

			// ---- DDA set-up (:309-351) ----
			float dirX = hd.x, dirY = hd.y;
			if (fabsf(dirY) < 1.e-4f)
				dirY = 1.e-4f;
			if (fabsf(dirX) < 1.e-4f)
				dirX = 1.e-4f;
			const float invDirX = fdiv(1.f, dirX), invDirY = fdiv(1.f, dirY);
			const bool posX = dirX > 0.f, posY = dirY > 0.f;
			dirBits = (posX ? 1u : 0u) | (posY ? 2u : 0u);
			invDirXI = (uint32_t)abs(osbn::x86_f2i(fmul(invDirX, 256.f)));
			invDirYI = (uint32_t)abs(osbn::x86_f2i(fmul(invDirY, 256.f)));
			dirXI = (uint32_t)abs(osbn::x86_f2i(fmul(dirX, 512.f)));
			dirYI = (uint32_t)abs(osbn::x86_f2i(fmul(dirY, 512.f)));

			const float cx = V.origin[0], cy = V.origin[1];
			cz = V.origin[2];
			icz = osbn::x86_f2i(floorf(cz));
			const int rx = osbn::x86_f2i(fmul(cx, 512.f)), ry = osbn::x86_f2i(fmul(cy, 512.f));
			irx = rx >> 9;
			iry = ry >> 9;
			const uint32_t fx = (uint32_t)rx & 511u, fy = (uint32_t)ry & 511u;
			gx = posX ? fx : 511u - fx;
			gy = posY ? fy : 511u - fy;

			distance = 1.e-20f;
			invDist = fdiv(1.f, distance);

			transScale = tn.y; // same expression as line.pitchScale (:303,358)
			transOffset = fmul(-tn.x, transScale);

			// zscale = Dot(horizonDir, viewAxis[2]) (horizonDir.z = 0), heightScale, depthBias (:403-414)
			zscale = fadd(fadd(fmul(hd.x, V.front[0]), fmul(hd.y, V.front[1])), fmul(0.f, V.front[2]));
			heightScale = V.front[2];
			depthBias = fmul(-cz, heightScale);

			tOld = tiled_xy((uint32_t)irx & 511u, (uint32_t)iry & 511u);
			count = 1;
			cnt2 = 1024; // static_cast<int>(fogDist * 8.f)
			advance(irx, iry, gx, gy, distance);
		}

		OSB_HDM void init(const WalkIO &io, uint32_t gl, uint32_t slot, uint32_t *filled, int fstride) {
			const int words = io.lineRes >> 5;
			for (int w = 0; w < words; w++)
				filled[w * fstride] = 0;
			init_consts(io, gl, slot);
			uint32_t m[4]; // parked behind the line's fill words (read once per step / per ceiling scan)
			eye_masks(m);
			for (int k = 0; k < 4; k++)
				filled[(words + k) * fstride] = m[k];
			sOld = io.solid[tOld];
			sNext = io.solid[nT];
		}

		// the line has ended: publish what it wrote
		OSB_HDM void finish(const WalkIO &io, const uint32_t *filled, int fstride) const {
			const int words = io.lineRes >> 5;
#if OSB_STRIP_HULL
			// hull [first, last] of the written strip pixels; the holes inside it (sky seen under an overhang, a bridge,
			// a floating block) get LinePixel::Clear, so that the gather reads the strip inside the hull and nothing else
			int first = 1, last = 0, w0 = 0, w1 = -1;
			for (int w = 0; w < words; w++) {
				const uint32_t m = filled[w * fstride];
				if (m) {
					if (w1 < 0) {
						w0 = w;
						first = (w << 5) + osbc::ffs32(m);
					}
					w1 = w;
					last = (w << 5) + 31 - osbc::clz32(m);
				}
			}
			const uint2 clearPix = make_uint2(0u, osbc::f2u(10000.f)); // LinePixel::Clear (:89-92)
			for (int w = w0; w <= w1; w++) {
				uint32_t holes = ~filled[w * fstride];
				if (w == w0)
					holes &= 0xffffffffu << (first & 31);
				if (w == w1)
					holes &= 0xffffffffu >> (31 - (last & 31));
				uint2 *dst = io.strips + (stripBase + ((uint32_t)w << 5));
				while (holes) {
					const int b = osbc::ffs32(holes);
					holes &= holes - 1;
					osbc::stcs(dst + b, clearPix);
				}
			}
			io.lineRecs[maskBase].view = first | ((last - first + 1) << 16);
#else
			uint32_t *out = io.lineMask + maskBase;
			for (int w = 0; w < words; w++)
				out[w] = filled[w * fstride];
#endif
		}

		// one candidate span: strip indices [from, to) get {colour of (cell, z) shaded by mode, depth of zA} where still empty
		OSB_HDM void fill(const WalkIO &io, uint32_t *filled, int fstride, int from, int to, uint32_t cell, int z, int zA,
		                  int mode, float medDist) const {
			using namespace osbc;
			const int w0 = from >> 5, w1 = (to - 1) >> 5;
			const uint32_t mFirst = 0xffffffffu << (from & 31), mLast = 0xffffffffu >> (31 - ((to - 1) & 31));
			uint2 px = make_uint2(0u, 0u);
			bool havePx = false;
			for (int w = w0; w <= w1; w++) {
				uint32_t m = 0xffffffffu;
				m = w == w0 ? m & mFirst : m;
				m = w == w1 ? m & mLast : m;
				const uint32_t curMask = filled[w * fstride];
				uint32_t todo = m & ~curMask;
				if (!todo)
					continue;
				filled[w * fstride] = curMask | todo;
				if (!havePx) {
					const uint32_t rawCol = ldg(io.color + (cell * 64u + (uint32_t)z));
					px = make_uint2(shade<kSse2Shade>(rawCol, mode), f2u(fadd(medDist, fmul(i2f(zA), heightScale))));
					havePx = true;
				}
				// (unsigned 32-bit pixel index: a chunk stays below 2^32 strip pixels, and the address is then one wide
				// multiply-add instead of a sign-extended 64-bit sum per store)
				const uint32_t wordBase = stripBase + ((uint32_t)w << 5);
				OSB_STAT(1, 1); // word with landing pixels
				do {
					const uint32_t b = (uint32_t)ffs32(todo);
					todo &= todo - 1;
					stcs(io.strips + (uint32_t)(wordBase + b), px);
					OSB_STAT(2, 1); // strip pixel stored
				} while (todo);
			}
		}

		// The spans of one DDA step in the reference's order: floor tops of the cell being left (ascending z, :584-603),
		// its ceilings (descending z, :604-622), then the walls of the cell being entered for the crossed face
		// (ascending z, :627-657).  Tops and walls share one loop (phase 0 / 1: both walk the lowest set bit first);
		// ceilings exist only with solid above the eye and run in between.  emit(from, to, cellIsTheEnteredOne, z, zA,
		// mode, ordinal) is called for every candidate whose strip interval [from, to) is not empty; ordinal counts the
		// step's candidates (empty ones included) in list order.  aboveEye(k) yields the lo / hi word of the z < icz mask.
		template <class AboveEye, class Emit>
		OSB_HDM void spans_of_step(const osbn::Num<kSse> &num, float vmax, uint64_t sOldV, uint64_t sNewV, float iDist, float oldIDist,
		                           int wallMode, uint32_t belowLo, uint32_t belowHi, AboveEye aboveEye, Emit emit) const {
			using namespace osbc;
			const uint32_t oLo = (uint32_t)sOldV, oHi = (uint32_t)(sOldV >> 32);
			const uint32_t nLo = (uint32_t)sNewV, nHi = (uint32_t)(sNewV >> 32);
			uint32_t cLo = oLo & ~(oLo << 1) & belowLo;                                 // solid, air above, below the eye
			uint32_t cHi = oHi & ~((oHi << 1) | (oLo >> 31)) & belowHi;
			const uint32_t wLo = nLo & ~oLo, wHi = nHi & ~oHi;                          // solid there, air in the cell left
			int phase = 0; // 0 tops, 1 walls
			int ordinal = 0;
			for (;;) {
				if (!(cLo | cHi)) {
					if (phase)
						break;
					// ceilings: solid, air below, above the eye; z = 63 is never listed (BuildRle :185-190)
					uint32_t bLo = oLo & ~((oLo >> 1) | (oHi << 31)) & aboveEye(0);
					uint32_t bHi = oHi & ~(oHi >> 1) & 0x7fffffffu & aboveEye(1);
					while (bLo | bHi) { // highest z first
						int z;
						if (bHi) {
							z = 63 - clz32(bHi);
							bHi &= ~(1u << (z - 32));
						} else {
							z = 31 - clz32(bLo);
							bLo &= ~(1u << z);
						}
						const int pA = transform(num, iDist, z + 1, vmax);
						const int pB = transform(num, oldIDist, z + 1, vmax);
						if (pB < pA)
							emit(pB, pA, false, z, z + 1, 2, ordinal);
						ordinal++;
					}
					phase = 1;
					cLo = wLo;
					cHi = wHi;
					if (!(cLo | cHi))
						break;
				}
				int z;
				if (cLo) {
					z = ffs32(cLo);
					cLo &= cLo - 1;
				} else {
					z = 32 + ffs32(cHi);
					cHi &= cHi - 1;
				}
				OSB_STAT(0, 1); // span candidate
				const int pA = transform(num, iDist, z, vmax);
				const int pB = transform(num, phase ? iDist : oldIDist, z + phase, vmax);
				if (pA < pB)
					emit(pA, pB, phase != 0, z, z, phase ? wallMode : 0, ordinal);
				ordinal++;
			}
		}

		// One DDA step with all its spans.  Returns false when the line has ended (its mask is published).
		OSB_HDM bool step(const WalkIO &io, const osbn::Num<kSse> &num, uint32_t *filled, int fstride) {
			using namespace osbc;
			const int lineRes = io.lineRes;
			const float vmax = fadd((float)lineRes, 0.5f);
			if (!(distance < kFogDist) || (--cnt2) == 0) { // while (distance < fogDist && --cnt2 > 0) (:423-425)
				finish(io, filled, fstride);
				return false;
			}
			// ---- take the precomputed step, start the one after it ----
			const uint64_t sNew = sNext;
			const uint32_t tNew = nT;
			const int wallMode = nWallMode;
			irx = nIrx;
			iry = nIry;
			gx = nGx;
			gy = nGy;
			distance = nDistance;
			advance(irx, iry, gx, gy, distance);
			sNext = io.solid[nT];

			const float oldInvDist = invDist;
			invDist = num.rcpHotPos(distance); // distance is in [1e-20, ~200]
			const float medDist = fadd(fmul(distance, zscale), depthBias);

			const int words = lineRes >> 5;
			spans_of_step(
			  num, vmax, sOld, sNew, invDist, oldInvDist, wallMode, filled[(words + 0) * fstride], filled[(words + 1) * fstride],
			  [&](int k) { return filled[(words + 2 + k) * fstride]; },
			  [&](int from, int to, bool cellNew, int z, int zA, int mode, int) {
				  fill(io, filled, fstride, from, to, cellNew ? tNew : tOld, z, zA, mode, medDist);
			  });
			sOld = sNew;
			tOld = tNew;
			// ---- pitch cull every 4th step (:661-667) ----
			if ((--count) == 0) {
				if ((transform(num, invDist, 0, vmax) >= lineRes - 1 && icz >= 0) || transform(num, invDist, 63, vmax) <= 0) {
					finish(io, filled, fstride);
					return false;
				}
				count = 4;
			}
			return true;
		}
	};

	// ------------------------------------------------------------------------------------------------------------
	// Cooperative line walker for SMALL launches (a single frame has only a few thousand yaw lines: one lane per line
	// leaves the GPU empty and a line's ~150 dependent steps set the frame time).  One WARP builds one line:
	//   path      one lane runs the map-independent DDA (:426-522) and records every step's cell and distance;
	//   batches   32 steps at a time, one lane per step: column masks, spans (spans_of_step, the same code the
	//             lane-per-line walker runs), the every-4th-step pitch cull (:661-667) by ballot;
	//   resolve   "first writer wins" along the strip is order, not time: a span's rank in the reference's sequence is
	//             (step, ordinal within the step).  Every non-empty span is recorded and does an atomicMin of
	//             rank << 11 | index on the strip pixels it covers, so a pixel ends up with exactly the span that would
	//             have written it first, whatever the order the lanes arrive in;
	//   store     one lane per strip pixel: winner -> colour (one load) + depth, stored as contiguous 256-byte runs;
	//             hull + LinePixel::Clear holes as LineWalker::finish does.
	// Same strips, bit for bit (tests/test_terrain_emul.py runs both walkers against the oracle).  A line with more
	// spans than kCoopMaxSpans falls back to the sequential walker.
	// ------------------------------------------------------------------------------------------------------------
	constexpr int kCoopMaxSteps = 1024; // the reference's own limit: cnt2 = fogDist * 8 (:424)
	constexpr int kCoopMaxSpans = 1024; // non-empty spans per line (measured: <= 160 at 1080p on Title.vxl)
	constexpr int kCoopShortSpan = 32;  // spans up to this many pixels are resolved by the lane that found them
	constexpr uint32_t kCoopEmpty = 0xffffffffu;

	struct CoopShared { // per warp; key[] (lineRes words) follows in dynamic shared memory
		uint32_t cell[kCoopMaxSteps]; // tiled cell entered by the step | wall mode << 31
		float dist[kCoopMaxSteps];    // distance after the step
		uint2 span[kCoopMaxSpans];    // x = from | to << 16, y = step << 17 | ordinal << 10 | z << 4 | mode << 2 | kind
		uint16_t longSpan[kCoopMaxSpans];
		int nSteps, nSpans, nLong;
	};
	OSB_HDI size_t coop_shared_bytes(int lineRes) { return sizeof(CoopShared) + sizeof(uint32_t) * (size_t)lineRes; }

	template <bool kSse, bool kSse2Shade> struct CoopLine {
		LineWalker<kSse, kSse2Shade> W;
		uint32_t tStart, eye[4];

		// every lane (uniform): constants of the line; the strip's pixel keys start empty
		OSB_HDM void begin(const WalkIO &io, uint32_t gl, uint32_t slot, CoopShared &sh, uint32_t *key, int lane) {
			W.init_consts(io, gl, slot);
			W.eye_masks(eye);
			tStart = W.tOld;
			for (int p = lane; p < io.lineRes; p += 32)
				key[p] = kCoopEmpty;
			if (lane == 0)
				sh.nSpans = sh.nLong = 0;
		}

		// one lane: the DDA of the whole line (it never looks at the map)
		OSB_HDM void path(CoopShared &sh) {
			int s = 0;
			while (W.distance < W.kFogDist && (--W.cnt2) != 0) { // while (distance < fogDist && --cnt2 > 0) (:423-425)
				sh.cell[s] = W.nT | ((uint32_t)W.nWallMode << 31);
				W.irx = W.nIrx;
				W.iry = W.nIry;
				W.gx = W.nGx;
				W.gy = W.nGy;
				W.distance = W.nDistance;
				sh.dist[s] = W.distance;
				W.advance(W.irx, W.iry, W.gx, W.gy, W.distance);
				s++;
			}
			sh.nSteps = s;
		}

		OSB_HDM float inv_dist_of(const osbn::Num<kSse> &num, const CoopShared &sh, int s) const {
			return s < 0 ? osbc::fdiv(1.f, 1.e-20f) : num.rcpHotPos(sh.dist[s]);
		}
		// does the line end after step s (pitch cull, checked after steps 0, 4, 8, ...)?
		OSB_HDM bool culls_after(const WalkIO &io, const osbn::Num<kSse> &num, const CoopShared &sh, int s) const {
			if (s >= sh.nSteps || (s & 3))
				return false;
			const float vmax = osbc::fadd((float)io.lineRes, 0.5f);
			const float iDist = inv_dist_of(num, sh, s);
			return (W.transform(num, iDist, 0, vmax) >= io.lineRes - 1 && W.icz >= 0) || W.transform(num, iDist, 63, vmax) <= 0;
		}

		OSB_HDM static uint32_t atomic_min(uint32_t *p, uint32_t v) {
#if defined(__CUDA_ARCH__)
			return atomicMin(p, v);
#else
			const uint32_t o = *p;
			if (v < o)
				*p = v;
			return o;
#endif
		}
		OSB_HDM static int atomic_inc(int *p) {
#if defined(__CUDA_ARCH__)
			return atomicAdd(p, 1);
#else
			return (*p)++;
#endif
		}

		// one lane per step: the spans of step s
		OSB_HDM void emit_step(const WalkIO &io, const osbn::Num<kSse> &num, CoopShared &sh, uint32_t *key, int s) {
			using namespace osbc;
			const float vmax = fadd((float)io.lineRes, 0.5f);
			const uint32_t tNew = sh.cell[s] & 0x7fffffffu, tOldS = s ? (sh.cell[s - 1] & 0x7fffffffu) : tStart;
			const int wallMode = (int)(sh.cell[s] >> 31);
			const uint64_t sOldV = io.solid[tOldS], sNewV = io.solid[tNew];
			const float iDist = inv_dist_of(num, sh, s), oldIDist = inv_dist_of(num, sh, s - 1);
			W.spans_of_step(
			  num, vmax, sOldV, sNewV, iDist, oldIDist, wallMode, eye[0], eye[1], [&](int k) { return eye[2 + k]; },
			  [&](int from, int to, bool cellNew, int z, int zA, int mode, int ordinal) {
				  const int idx = atomic_inc(&sh.nSpans);
				  if (idx >= kCoopMaxSpans)
					  return; // overflow: the caller falls back to the sequential walker
				  const uint32_t kind = cellNew ? 2u : (zA != z ? 1u : 0u);
				  sh.span[idx] = make_uint2((uint32_t)from | ((uint32_t)to << 16),
				                            ((uint32_t)s << 17) | ((uint32_t)ordinal << 10) | ((uint32_t)z << 4) | ((uint32_t)mode << 2) | kind);
				  if (to - from <= kCoopShortSpan) {
					  const uint32_t k = ((((uint32_t)s << 7) | (uint32_t)ordinal) << 11) | (uint32_t)idx;
					  for (int p = from; p < to; p++)
						  atomic_min(key + p, k);
				  } else {
					  sh.longSpan[atomic_inc(&sh.nLong)] = (uint16_t)idx;
				  }
			  });
		}
		// all lanes: the long spans, 32 pixels at a time
		OSB_HDM void resolve_long(const CoopShared &sh, uint32_t *key, int lane) const {
			for (int i = 0; i < sh.nLong; i++) {
				const int idx = sh.longSpan[i];
				const uint2 r = sh.span[idx];
				const int from = (int)(r.x & 0xffffu), to = (int)(r.x >> 16);
				const uint32_t k = ((r.y >> 10) << 11) | (uint32_t)idx; // (step << 7 | ordinal) << 11 | index
				for (int p = from + lane; p < to; p += 32)
					atomic_min(key + p, k);
			}
		}
		// this lane's share of the hull: lowest / highest written strip pixel among p = lane, lane + 32, ...
		OSB_HDM void scan_hull(const WalkIO &io, const uint32_t *key, int lane, int &first, int &last) const {
			first = io.lineRes;
			last = -1;
			for (int p = lane; p < io.lineRes; p += 32)
				if (key[p] != kCoopEmpty) {
					first = osbc::imin(first, p);
					last = osbc::imax(last, p);
				}
		}
		// one lane per strip pixel inside the hull [first, last]; kB pixels per lane and pass, stage by stage, so that
		// the colour loads of a pass (DRAM latency on a cold map) are in flight together
		OSB_HDM void store(const WalkIO &io, const CoopShared &sh, const uint32_t *key, int lane, int first, int last) const {
			using namespace osbc;
			constexpr int kB = 4;
			const uint2 clearPix = make_uint2(0u, f2u(10000.f)); // LinePixel::Clear (:89-92)
			for (int p0 = (first & ~31) + lane; p0 <= last; p0 += 32 * kB) {
				uint32_t spanY[kB], rawCol[kB];
				bool live[kB];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
				for (int j = 0; j < kB; j++) {
					const int p = p0 + 32 * j;
					live[j] = false;
					spanY[j] = 0u;
					rawCol[j] = 0u;
					if (p >= first && p <= last) {
						const uint32_t k = key[p];
						if (k != kCoopEmpty) {
							live[j] = true;
							spanY[j] = sh.span[k & (uint32_t)(kCoopMaxSpans - 1)].y;
						}
					}
				}
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
				for (int j = 0; j < kB; j++) {
					if (live[j]) {
						const int s = (int)(spanY[j] >> 17), z = (int)((spanY[j] >> 4) & 63u);
						const uint32_t kind = spanY[j] & 3u;
						const uint32_t cellIdx = kind == 2u ? (sh.cell[s] & 0x7fffffffu) : (s ? (sh.cell[s - 1] & 0x7fffffffu) : tStart);
						rawCol[j] = ldg(io.color + (cellIdx * 64u + (uint32_t)z));
					}
				}
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
				for (int j = 0; j < kB; j++) {
					const int p = p0 + 32 * j;
					if (p < first || p > last)
						continue;
					uint2 px = clearPix; // a hole inside the hull
					if (live[j]) {
						const int s = (int)(spanY[j] >> 17), z = (int)((spanY[j] >> 4) & 63u), mode = (int)((spanY[j] >> 2) & 3u);
						const uint32_t kind = spanY[j] & 3u;
						const float medDist = fadd(fmul(sh.dist[s], W.zscale), W.depthBias);
						px = make_uint2(shade<kSse2Shade>(rawCol[j], mode),
						                f2u(fadd(medDist, fmul(i2f(z + (kind == 1u ? 1 : 0)), W.heightScale))));
					}
					stcs(io.strips + (W.stripBase + (uint32_t)p), px);
				}
			}
			if (lane == 0) {
#if OSB_STRIP_HULL
				io.lineRecs[W.maskBase].view = last >= first ? (first | ((last - first + 1) << 16)) : 1;
#endif
			}
		}
	};

	// ---- fastATan / fastATan2 (SWMapRenderer.cpp:697-742); tab = sm | lg | smN | lgN, 5000 each ----
	template <bool kSse> OSB_HDI uint32_t fast_atan(const osbn::Num<kSse> &num, const uint16_t *__restrict__ tab, float v) {
		using namespace osbc;
		int idx, base;
		if (v < 0.f) {
			if (v > -1.f) {
				idx = f2i_rz(fmul(v, -4096.f));
				base = 10000;
			} else {
				idx = f2i_rz(num.div(-4096.f, v));
				base = 15000;
			}
		} else {
			if (v < 1.f) {
				idx = f2i_rz(fmul(v, 4096.f));
				base = 0;
			} else {
				idx = f2i_rz(num.div(4096.f, v));
				base = 5000;
			}
		}
		// |v| < 1 -> |v| * 4096 < 4096, |v| >= 1 -> 4096 / |v| <= 4096: cvttss2si's out-of-range case cannot occur, and a NaN
		// (index "indefinite" in the reference, an out-of-bounds read there) ends as entry 0 here
		idx = imin(imax(idx, 0), 4999); // memory safety only; in range for every finite input
		return ldg(tab + base + idx);
	}
	template <bool kSse>
	OSB_HDI uint32_t fast_atan2(const osbn::Num<kSse> &num, const uint16_t *__restrict__ tab, float y, float x) {
		if (x == 0.f)
			return y > 0.f ? 16384u : 49152u; // (uint16_t)-16384
		else if (x > 0.f)
			return fast_atan(num, tab, num.div(y, x));
		return (fast_atan(num, tab, num.div(y, x)) + 32768u) & 0xffffu;
	}

	// ApplyFog blend of one pixel (SWRenderer.cpp:381-407 scalar, :462-510 SSE2).  col = B | G<<8 | R<<16 | X<<24.
	// B and R share one multiply: each 8-bit channel times a factor <= 256 stays below 2^16.
	template <bool kSse2Fog> OSB_HDI uint32_t fog_blend(uint32_t col, float dist, uint32_t fogBR, uint32_t fogG8) {
		int factor;
		if (kSse2Fog) {
			float d = dist > 0.f ? dist : 0.f; // maxps / minps return the 2nd operand on NaN
			d = d < 256.f ? d : 256.f;
			factor = osbc::f2i_rn(d); // d is in [0, 256]: cvtps2dq's out-of-range case cannot occur
		} else {
			factor = osbc::imin(osbn::x86_f2i(dist), 256);
			factor = osbc::imax(0, factor);
		}
		const uint32_t f1 = (uint32_t)factor, f2 = (uint32_t)(256 - factor);
		const uint32_t br = (col & 0x00ff00ffu) * f2 + fogBR * f1; // B' and R' are bytes 1 and 3
		// G and X (bytes 1 and 3) share the second multiply like B and R the first: low half G * f2 + fogG * f1, high half
		// X * f2, both below 2^16; their bits 8..15 are the result bytes and already sit where they belong
		const uint32_t gx = ((col >> 8) & 0x00ff00ffu) * f2 + (fogG8 >> 8) * f1;
		// lane 3 of the SSE2 path keeps (X * f2) >> 8, the scalar path writes 0 there
		// result bytes: B' R' = bytes 1 and 3 of the first sum, G' X' = bytes 1 and 3 of the second: one byte permute picks
		// all four (instead of a shift, two masks and an or)
		const uint32_t out = osbc::byte_perm(br, gx, 0x7351);
		return kSse2Fog ? out : (out & 0x00ffffffu);
	}

	// ------------------------------------------------------------------------------------------------------------
	// RenderFinal + ApplyFog (SWMapRenderer.cpp:744-1043, SWRenderer.cpp:341-516).  One warp per (view, 32-pixel
	// column group = four 8x8 block columns, segment of kSegRows block rows).  Every lane has two roles:
	//   corner role   lane = r*16 + b*4 + k evaluates corner k of block column b in block row by + r -- the 32 block
	//                 corners of TWO block rows per pass (atan-LUT yaw, rsqrt pitch), published through GatherShared;
	//   pixel role    lane = b*8 + px owns one pixel column: the reference's integer interpolation between the four
	//                 corners of its block, the gather from (line, pitch index) and the fog blend, written once:
	//                 every pixel row is one 128-byte run per buffer.
	// ------------------------------------------------------------------------------------------------------------
	constexpr int kGatherWarps = 4;
	constexpr int kSegRows = 8; // most 8x8 block rows per warp task (vertical); GatherIO::segRows picks the actual number

	struct GatherShared { // per warp
		int yaw[32], pit[32]; // [r*16 + b*4 + k]
		float v2[2][4][3];    // RenderFinal's v2 of block (b, by + r)
		float fog[kSegRows * 2][8];
	};

	struct GatherIO {
		const ViewParams *views;
		int nViews;
		FrameGeom g;
		const float *tabs;
		const uint64_t *solid;
		const uint16_t *atanTab;
		const uint2 *strips;
		const LineRec *lineRecs;
		const float2 *lineTan;
		const uint32_t *lineMask;
		int lineRes;
		uint32_t *fb;
		float *depth;
		int segRows; // block rows per warp task, 1..kSegRows (small launches use fewer: more warps in flight)
	};

	// kCg: strips and line records are read with ld.global.cg (L2 only).  Needed when the SAME kernel wrote them on another
	// SM (k_view_pipeline): L1 is not coherent, and this SM's L1 may still hold a record line its walker warps read before
	// the line's hull was published.
	template <bool kSse, bool kSse2Fog, bool kCg = false> struct GatherLane {
		// task
		int by0, by1, colBlocks, cg;
		const ViewParams *V;
		uint32_t *fbLane;      // frame pixel (row 0 of the task's first block row, this lane's column) of this lane's view
		float *dbLane;         // ... and its depth pixel: an address is then one 32-bit offset away (one wide multiply-add)
		const LineRec *lpLane; // line records of the view
		// pixel role
		int pb, px, xs, hSh;
		uint32_t o; // pixel offset inside the view of this lane's column at block row by0
		bool colValid, skip;
		// corner role
		int cr, ck, cb;
		float v2c[3], dRL[3], dDL[3];
		// line records of this lane's view staged in shared memory by the kernel (bulk copy), or null: read them from global
		const LineRec *stagedRecs = nullptr;

		// false: the task lies outside the launch (whole warp)
		OSB_HDM bool init(const GatherIO &io, uint32_t task, int lane) {
			using namespace osbc;
			const FrameGeom &g = io.g;
			colBlocks = g.fw >> 3;
			const int rowBlocks = g.fh >> 3;
			const int colGroups = (colBlocks + 3) >> 2;
			const int segs = (rowBlocks + io.segRows - 1) / io.segRows;
			const uint32_t perView = (uint32_t)colGroups * segs;
			if (task >= perView * (uint32_t)io.nViews)
				return false;
			const int vi = task / perView;
			const uint32_t rem = task - (uint32_t)vi * perView;
			const int seg = rem / colGroups;
			cg = rem - seg * colGroups;
			V = io.views + vi;
			by0 = seg * io.segRows;
			by1 = imin(by0 + io.segRows, rowBlocks);

			pb = lane >> 3;
			const int bx = (cg << 2) + pb;     // this lane's block column
			colValid = bx < colBlocks;         // a partial last group when fw is not a multiple of 32
			const int bxc = colValid ? bx : colBlocks - 1;
			px = lane & 7;                     // pixel column inside the block
			o = (uint32_t)(by0 * 8) * (uint32_t)g.fw + (uint32_t)((bxc << 3) + px);
			fbLane = io.fb + (size_t)vi * g.fw * g.fh + o;
			dbLane = io.depth + (size_t)vi * g.fw * g.fh + o;
			lpLane = io.lineRecs + V->lineBase;
			const int under = g.under;
			hSh = under == 1 ? 3 : (under == 2 ? 2 : 1); // log2(8 / under)
			xs = (px - (px % under)) / under;            // undersampling: pixels share their group leader's sample

			// SWMapRenderer::Render :1181-1184: eye inside a solid voxel (or no map at all) -> terrain skipped, fb keeps
			// the clear colour and the depth buffer keeps its previous contents
			{
				const int ex = osbn::x86_f2i(floorf(V->origin[0])), ey = osbn::x86_f2i(floorf(V->origin[1])),
				          ez = osbn::x86_f2i(floorf(V->origin[2]));
				skip = g.noMap || ez >= 64 ||
				       (ez >= 0 && ((io.solid[tiled_xy((uint32_t)ex & 511u, (uint32_t)ey & 511u)] >> ez) & 1ull));
			}

			// corner role
			cr = lane >> 4;
			cb = (lane >> 2) & 3;
			ck = lane & 3;
			const int cbx = imin((cg << 2) + cb, colBlocks - 1);
			const float *v1Tab = io.tabs + V->tabBase + (g.fw >> 2) + (g.fh >> 2);
			const float sD = fdiv(fmul(V->fovYt, 2.f), (float)g.fh);
			const float sR = fmul(fdiv(fmul(V->fovXt, 2.f), (float)g.fw), (float)under);
			for (int k = 0; k < 3; k++) { // per-view deltas (:757-775)
				dDL[k] = fmul(fmul(-V->up[k], sD), 8.f);
				dRL[k] = fmul(fmul(V->right[k], sR), (float)(8 / under));
				v2c[k] = v1Tab[cbx * 3 + k];
			}
			for (int i = 0; i < by0 + cr; i++) { // replay v2 += deltaDownLarge (:1037)
				for (int k = 0; k < 3; k++)
					v2c[k] = fadd(v2c[k], dDL[k]);
			}
			return true;
		}

		// fog scale of every 4x4 pixel block this task covers (SWRenderer.cpp:372-380); lane-strided over the task
		OSB_HDM void fog_phase(const GatherIO &io, const osbn::Num<kSse> &num, GatherShared &sm, int lane) const {
			using namespace osbc;
			const FrameGeom &g = io.g;
			const float *vxTab = io.tabs + V->tabBase, *vyTab = vxTab + (g.fw >> 2);
			const int nRows4 = (by1 - by0) * 2;
			for (int i = lane; i < nRows4 * 8; i += 32) {
				const int r4 = i >> 3, c4 = i & 7;
				const int cx4 = imin((cg << 3) + c4, (g.fw >> 2) - 1);
				const float vx = vxTab[cx4], vy = vyTab[(by0 << 1) + r4];
				float ds = fadd(fadd(1.f, fmul(vx, vx)), fmul(vy, vy));
				ds = fmul(ds, fmul(num.rsqrt(ds), V->fogScale));
				sm.fog[r4][c4] = ds;
			}
		}

		// block corners of block rows by, by + 1 (:810-835), then this lane's corner vector moves two rows down
		OSB_HDM void corner_phase(const GatherIO &io, const osbn::Num<kSse> &num, GatherShared &sm, int lane, int by) {
			using namespace osbc;
			int yaw = 0, pit = 0;
			if (by + cr < by1) {
				const bool slow = v2c[2] > 0.99f || v2c[2] < -0.99f;
				if (ck == 0) {
					sm.v2[cr][cb][0] = v2c[0];
					sm.v2[cr][cb][1] = v2c[1];
					sm.v2[cr][cb][2] = v2c[2];
				}
				if (!slow && !skip) {
					float c[3];
					for (int k = 0; k < 3; k++) {
						c[k] = v2c[k];
						if (ck & 1)
							c[k] = fadd(c[k], dRL[k]);
						if (ck & 2)
							c[k] = fadd(c[k], dDL[k]);
					}
					yaw = (int)((((int)fast_atan2(num, io.atanTab, c[1], c[0]) - V->yawMin2) & 0xffff) << 8);
					float pitch = fmul(c[2], num.rsqrt(fadd(fmul(c[0], c[0]), fmul(c[1], c[1]))));
					pitch = num.toSpecialTan(pitch);
					pit = osbn::x86_f2i(fmul(pitch, 65536.f * 8192.f));
				}
			}
			sm.yaw[lane] = yaw;
			sm.pit[lane] = pit;
			for (int k = 0; k < 3; k++)
				v2c[k] = fadd(fadd(v2c[k], dDL[k]), dDL[k]);
		}

#if OSB_STRIP_HULL
		// LinePixel of (line record, strip index): one predicated load inside the line's hull, LinePixel::Clear outside
		OSB_HDM uint32_t fetch_mask(const GatherIO &, int, int) const { return 0u; }
		OSB_HDM uint2 fetch_pixel(const GatherIO &io, int slot, int pitchIndex, uint32_t hull) const {
			uint2 pix = make_uint2(0u, osbc::f2u(10000.f)); // LinePixel::Clear (:89-92)
			// first <= pitchIndex < first + count as one unsigned compare; a culled line carries count = 0
			if ((uint32_t)(pitchIndex - (int)(hull & 0xffffu)) < (hull >> 16)) {
				const uint2 *p = io.strips + ((uint32_t)slot + (uint32_t)pitchIndex); // slot = first strip pixel of the line
				pix = kCg ? osbc::ldcg(p) : osbc::ldg(p);
			}
			return pix;
		}
#else
		// LinePixel of (strip slot, strip index pitchIndex): two predicated loads, no nested branch, so that the loads of
		// the pixels of a batch can be in flight together.  mask word first: unwritten strip memory is never read.
		OSB_HDM uint32_t fetch_mask(const GatherIO &io, int slot, int pitchIndex) const {
			uint32_t mw = 0u; // culled line: every strip pixel is LinePixel::Clear
			if (slot >= 0)
				mw = osbc::ldg(io.lineMask + ((uint32_t)slot * (uint32_t)(io.lineRes >> 5) + ((uint32_t)pitchIndex >> 5)));
			return mw;
		}
		OSB_HDM uint2 fetch_pixel(const GatherIO &io, int slot, int pitchIndex, uint32_t mw) const {
			uint2 pix = make_uint2(0u, osbc::f2u(10000.f)); // LinePixel::Clear (:89-92)
			if ((mw >> (pitchIndex & 31)) & 1u)
				pix = osbc::ldg(io.strips + ((uint32_t)slot * (uint32_t)io.lineRes + (uint32_t)pitchIndex));
			return pix;
		}
#endif

		// the 8 pixel rows of block row by + r for this lane's pixel column
		OSB_HDM void pixel_phase(const GatherIO &io, const osbn::Num<kSse> &num, const GatherShared &sm, int r, int by) {
			using namespace osbc;
			const FrameGeom &g = io.g;
			const int row = by + r;
			if (row >= by1)
				return;
			// one 64-bit pixel pointer per lane; the depth pixel lies a constant distance behind it
			uint32_t *fbV = fbLane;
			float *dbV = dbLane;
			uint32_t oo = (uint32_t)((row - by0) * 8) * (uint32_t)g.fw;
			const float fog0 = sm.fog[(row - by0) * 2][px >> 2 | (pb << 1)], fog1 = sm.fog[(row - by0) * 2 + 1][px >> 2 | (pb << 1)];
			const uint32_t fogBR = g.fogB | (g.fogR << 16), fogG8 = g.fogG << 8;
			if (skip) {
				if (colValid) {
					for (int py = 0; py < 8; py++, oo += g.fw) {
						const float dep = dbV[oo];
						fbV[oo] = g.noFog ? 0x7f7f7fu : fog_blend<kSse2Fog>(0x7f7f7fu, fmul(dep, py < 4 ? fog0 : fog1), fogBR, fogG8);
					}
				}
				return;
			}
			if (!colValid)
				return;
			const float v2z = sm.v2[r][pb][2];
			const bool slow = v2z > 0.99f || v2z < -0.99f; // uniform within a block (8 lanes)
			const uint32_t numLines = (uint32_t)V->numLines;
			const uint32_t yawScale2 = (uint32_t)V->yawScale2;
			const int lineRes = io.lineRes;
			const LineRec *lpV = lpLane;
			if (!slow) {
				const int ci = r * 16 + pb * 4;
				const int yaw1 = sm.yaw[ci], yaw2 = sm.yaw[ci + 1], yaw3 = sm.yaw[ci + 2], yaw4 = sm.yaw[ci + 3];
				const int pit1 = sm.pit[ci], pit2 = sm.pit[ci + 1], pit3 = sm.pit[ci + 2], pit4 = sm.pit[ci + 3];
				// bilinear interpolation in integers; `<<8>>8` is the reference's phase unwrap (:837-881)
				const int yawDiff1 = tdiv_pow2(((yaw2 - yaw1) << 8) >> 8, hSh), yawDiff2 = tdiv_pow2(((yaw4 - yaw3) << 8) >> 8, hSh);
				const int pitDiff1 = tdiv_pow2(pit2 - pit1, hSh), pitDiff2 = tdiv_pow2(pit4 - pit3, hSh);
				const int yawA = yaw1 + xs * yawDiff1, yawB = yaw3 + xs * yawDiff2;
				const int pitA = pit1 + xs * pitDiff1, pitB = pit3 + xs * pitDiff2;
				const int dYaw = tdiv_pow2(((yawB - yawA) << 8) >> 8, 3), dPit = tdiv_pow2(pitB - pitA, 3);
				int yawC = yawA, pitC = pitA;
				constexpr int kB = OSB_GATHER_BATCH;
				// kB (four) pixel rows per batch, stage by stage: the four line records, then the four mask words, then the
				// four strip pixels are requested together (each stage depends on the one before it)
				// (the two halves and the two block rows of a pass share ONE copy of this code: fully unrolled the kernel
				// was 4096 instructions and stalled 6 cycles per issue on instruction fetch)
#if defined(__CUDA_ARCH__)
				OSB_PRAGMA(unroll OSB_GATHER_HALF_UNROLL)
#endif
				for (int half = 0; half < 8 / kB; half++) {
					int slot[kB], pitchIndex[kB];
					int4 lpw[kB];
					uint32_t mw[kB];
					uint2 pix[kB];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
					for (int j = 0; j < kB; j++, yawC += dYaw) {
						uint32_t yawIndex = (uint32_t)((yawC << 8) >> 16);
						yawIndex = (yawIndex * yawScale2) >> 16;
						yawIndex = (yawIndex * numLines) >> 16;
						yawIndex = uminu(yawIndex, numLines - 1); // memory safety; the reference would read a stale line
						lpw[j] = stagedRecs ? *reinterpret_cast<const int4 *>(stagedRecs + yawIndex)
						                    : (kCg ? ldcg(reinterpret_cast<const int4 *>(lpV + yawIndex))
						                           : ldg(reinterpret_cast<const int4 *>(lpV + yawIndex)));
					}
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
					for (int j = 0; j < kB; j++, pitC += dPit) {
						int pi = (pitC >> 13) - lpw[j].x; // LineRec::pitchTanMinI
						pi = mulhi(pi, lpw[j].y);         // LineRec::pitchScaleI
						pitchIndex[j] = pi & (lineRes - 1);
						slot[j] = lpw[j].z;
#if OSB_STRIP_HULL
						mw[j] = (uint32_t)lpw[j].w;
#else
						mw[j] = fetch_mask(io, slot[j], pitchIndex[j]);
#endif
					}
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
					for (int j = 0; j < kB; j++)
						pix[j] = fetch_pixel(io, slot[j], pitchIndex[j], mw[j]);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
					for (int j = 0; j < kB; j++, oo += g.fw) {
						const float dep = u2f(pix[j].y);
						stcs(fbV + oo, g.noFog ? pix[j].x : fog_blend<kSse2Fog>(pix[j].x, fmul(dep, (half * kB + j) < 4 ? fog0 : fog1), fogBR, fogG8));
						stcs(dbV + oo, dep);
					}
				}
			} else {
				// exact per-pixel path near the poles (:940-1033)
				const float sD = fdiv(fmul(V->fovYt, 2.f), (float)g.fh);
				const float sR = fmul(fdiv(fmul(V->fovXt, 2.f), (float)g.fw), (float)g.under);
				float v4[3], dD[3], dR[3];
				for (int k = 0; k < 3; k++) {
					dD[k] = fmul(-V->up[k], sD);
					dR[k] = fmul(V->right[k], sR);
					v4[k] = sm.v2[r][pb][k];
				}
				for (int i = 0; i < xs; i++) {
					for (int k = 0; k < 3; k++)
						v4[k] = fadd(v4[k], dR[k]);
				}
				const float2 *ltV = io.lineTan + V->lineBase;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
				for (int py = 0; py < 8; py++, oo += g.fw) {
					uint32_t yawIndex = (uint32_t)(((int)fast_atan2(num, io.atanTab, v4[1], v4[0]) - V->yawMin2) & 0xffff);
					yawIndex = (yawIndex * yawScale2) >> 16;
					yawIndex = (yawIndex * numLines) >> 16;
					yawIndex = osbc::uminu(yawIndex, numLines - 1);
					LineRec lp;
					if (kCg) {
						const int4 w4 = ldcg(reinterpret_cast<const int4 *>(lpV + yawIndex));
						lp.pitchTanMinI = w4.x;
						lp.pitchScaleI = w4.y;
						lp.slot = w4.z;
						lp.view = w4.w;
					} else {
						lp = lpV[yawIndex];
					}
					const float2 lt = ltV[yawIndex];
					float pitch = fmul(v4[2], num.rsqrt(fadd(fmul(v4[0], v4[0]), fmul(v4[1], v4[1]))));
					pitch = num.toSpecialTan(pitch);
					pitch = fmul(fsub(pitch, lt.x), lt.y);
					const int pitchIndex = osbn::x86_f2i(pitch) & (lineRes - 1);
#if OSB_STRIP_HULL
					const uint2 pix = fetch_pixel(io, lp.slot, pitchIndex, (uint32_t)lp.view);
#else
					const uint2 pix = fetch_pixel(io, lp.slot, pitchIndex, fetch_mask(io, lp.slot, pitchIndex));
#endif
					const float dep = u2f(pix.y);
					stcs(fbV + oo, g.noFog ? pix.x : fog_blend<kSse2Fog>(pix.x, fmul(dep, py < 4 ? fog0 : fog1), fogBR, fogG8));
					stcs(dbV + oo, dep);
					for (int k = 0; k < 3; k++)
						v4[k] = fadd(v4[k], dD[k]);
				}
			}
		}
	};

} // namespace osb
