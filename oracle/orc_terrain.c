/*
 * TEST INFRASTRUCTURE ONLY (see osb_oracle.h) -- CPU restatement of the reference's software
 * terrain pass and fog for one view:
 *   SWRenderer::EndScene            Draw/SWRenderer.cpp:709-745   (clear, terrain, fog)
 *   SWRenderer::BuildFrustrum       Draw/SWRenderer.cpp:207-227
 *   SWMapRenderer::Render/RenderInner  Draw/SWMapRenderer.cpp:1045-1194
 *   SWMapRenderer::BuildLine        Draw/SWMapRenderer.cpp:219-671
 *   SWMapRenderer::RenderFinal      Draw/SWMapRenderer.cpp:744-1043
 *   SWRenderer::ApplyFog            Draw/SWRenderer.cpp:341-516
 * in the three numerics variants the reference has (Draw/SWFeatureLevel.h:25-121).
 *
 * Sequential, single-threaded; the reference's thread count only matters through the per-thread
 * float recurrences of RenderFinal / ApplyFog, which are replayed here per emulated thread.
 * Compile with -ffp-contract=off and no -march flags.
 */
#include "osb_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <xmmintrin.h>

#define MW 512
#define MH 512

typedef struct {
	float x, y, z;
} v3;
static inline v3 V3(float x, float y, float z) {
	v3 r = {x, y, z};
	return r;
}
static inline v3 v3add(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 v3sub(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 v3mul(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
static inline float v3dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; } /* Core/Math.h:347-349 */
static inline v3 v3norm(v3 a) {                                                     /* Core/Math.h:369-374 */
	float s = sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
	if (s != 0.f)
		s = 1.f / s;
	return V3(a.x * s, a.y * s, a.z * s);
}

/* ---- numerics variants: Draw/SWFeatureLevel.h:89-121 ---- */
static int g_sse; /* 0 = IEEE build, 1 = rcpss/rsqrtss build */
static inline float hw_rcp(float b) { return _mm_cvtss_f32(_mm_rcp_ss(_mm_set_ss(b))); }
static inline float hw_rsqrt(float b) { return _mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(b))); }
static inline float fast_rcp(float b) { return g_sse ? hw_rcp(b) : 1.f / b; }
static inline float fast_div(float a, float b) { return g_sse ? a * hw_rcp(b) : a / b; }
static inline float fast_rsqrt(float b) { return g_sse ? hw_rsqrt(b) : 1.f / sqrtf(b); }
static inline float fast_sqrt(float s) {
	if (s == 0.f)
		return 0.f;
	return s * fast_rsqrt(s);
}

/* SWMapRenderer.cpp:43-66 */
static float special_tan(float v) {
	static const float pi = M_PI;
	if (v <= -pi * 0.5f)
		return -2.f;
	else if (v < -pi * 0.25f)
		v = -2.f - 1.f / tanf(v);
	else if (v < pi * 0.25f)
		v = tanf(v);
	else if (v < pi * 0.5f)
		v = 2.f - 1.f / tanf(v);
	else
		return 2.f;
	return v;
}
static inline float to_special_tan(float v) {
	if (v < -1.f)
		return -2.f - fast_rcp(v);
	else if (v > 1.f)
		return 2.f - fast_rcp(v);
	return v;
}

enum { F_POSX = 0, F_NEGX = 1, F_POSY = 2, F_NEGY = 3, F_POSZ = 4, F_NEGZ = 5 }; /* SWMapRenderer.cpp:68 */

typedef struct {
	uint32_t combined;
	float depth;
} line_px;

typedef struct {
	v3 horizonDir;
	float pitchTanMin, pitchScale;
	int pitchTanMinI, pitchScaleI;
} line_t;

typedef struct {
	const uint64_t *solid;
	const uint32_t *color;
	orc_scene sc;
	v3 origin, axis[3];
	v3 frustumN[6];
	int numerics;
	int lineRes;
	int numLines;
	line_t *lines;
	line_px *pixels; /* numLines * lineRes */
} frame_t;

static inline uint64_t solid_w(const frame_t *f, int x, int y) {
	return f->solid[(size_t)(x & (MW - 1)) * MH + (y & (MH - 1))];
}
static inline uint32_t color_w(const frame_t *f, int x, int y, int z) {
	return f->color[((size_t)(x & (MW - 1)) * MH + (y & (MH - 1))) * 64 + (z & 63)];
}

/* BuildLinePixel, SWMapRenderer.cpp:532-576 */
static line_px build_line_pixel(const frame_t *f, int x, int y, int z, int face, float dist) {
	line_px px;
	px.depth = dist;
	uint32_t col = color_w(f, x, y, z);
	if (f->numerics == ORC_NUM_SSE2) {
		/* SSE2 path :536-557: 16-bit lanes {R,G,B} swapped to B,G,R order, shifts, saturating pack */
		uint32_t ch[3] = {(col >> 16) & 0xff, (col >> 8) & 0xff, col & 0xff}; /* byte0=B byte1=G byte2=R of px */
		for (int k = 0; k < 3; k++) {
			uint32_t m = ch[k];
			switch (face) {
				case F_POSZ: m >>= 1; break;
				case F_POSX:
				case F_POSY:
				case F_NEGX: m = (m >> 1) + (m >> 2); break;
				default: break;
			}
			if ((col >> 24) < 100)
				m >>= 1;
			ch[k] = m > 255 ? 255 : m;
		}
		/* lane 3 carries the health byte through the same shifts; it is overwritten by `filled` below
		   only in bit 0 of byte 3 -- reproduce the byte exactly */
		uint32_t a = (col >> 24) & 0xff;
		switch (face) {
			case F_POSZ: a >>= 1; break;
			case F_POSX:
			case F_POSY:
			case F_NEGX: a = (a >> 1) + (a >> 2); break;
			default: break;
		}
		if ((col >> 24) < 100)
			a >>= 1;
		if (a > 255)
			a = 255;
		px.combined = ch[0] | (ch[1] << 8) | (ch[2] << 16) | (a << 24);
		px.combined |= 1u << 24; /* px.filled = true (bit 24, :79) */
	} else {
		col = (col & 0xff00) | ((col & 0xff) << 16) | ((col & 0xff0000) >> 16);
		switch (face) {
			case F_POSZ: col = (col & 0xfcfcfc) >> 2; break;
			case F_POSX:
			case F_POSY:
			case F_NEGX: col = (col & 0xfefefe) >> 1; break;
			default: break;
		}
		px.combined = col | (1u << 24);
	}
	return px;
}

/* pitch clip of one frustum side plane, SWMapRenderer.cpp:243-283 */
static void clip_plane(v3 plane, v3 horz, float *minPitch, float *maxPitch) {
	if (plane.x == 0.f && plane.y == 0.f) {
		if (plane.z > 0.f)
			*minPitch = *minPitch < 0.f ? 0.f : *minPitch; /* std::max(minPitch, 0.f) */
		else
			*maxPitch = 0.f < *maxPitch ? 0.f : *maxPitch; /* std::min(maxPitch, 0.f) */
	} else if (plane.z == 0.f) {
		if (v3dot(plane, horz) < 0.f) {
			*minPitch = 2.f;
			*maxPitch = -2.f;
		}
	} else {
		v3 prj = plane;
		prj.z = 0.f;
		prj = v3norm(prj);
		float zv = fabsf(plane.z);
		float cs = v3dot(prj, horz);
		float ang = zv * zv * (1.f - cs * cs) / (cs * cs);
		ang = -cs * fast_sqrt(1.f + ang);
		ang = zv / ang;
		if (isnan(ang) || isinf(ang) || ang == 0.f)
			return;
		ang = fast_sqrt(1.f - ang * ang) / ang;
		ang = atanf(ang);
		if (isnan(ang) || isinf(ang))
			return;
		if (plane.z > 0.f) {
			float c = ang - 0.01f;
			*minPitch = *minPitch < c ? c : *minPitch; /* std::max(a,b) = (a<b)?b:a */
		} else {
			float c = -ang + 0.01f;
			*maxPitch = c < *maxPitch ? c : *maxPitch; /* std::min(a,b) = (b<a)?b:a */
		}
	}
}

typedef struct {
	float zval[65];
	float transScale, transOffset, vmax;
} xform_t;

/* transform lambda, SWMapRenderer.cpp:395-401 */
static inline int transform(const xform_t *t, float invDist, int z) {
	float p = to_special_tan(invDist * t->zval[z]) * t->transScale + t->transOffset;
	p = p < 0.f ? 0.f : p;         /* std::max(p, 0.f): NaN stays NaN -> (a<b)?b:a with a=p */
	p = t->vmax < p ? t->vmax : p; /* std::min(p, vmax) */
	return (int)p; /* static_cast<uint_fast16_t>; 0 <= p <= lineResolution + 0.5 */
}

static inline void fill_span(line_px *pixels, int from, int to, line_px pix) {
	for (int j = from; j < to; j++)
		if (pixels[j].combined == 0)
			pixels[j] = pix;
}

/* SWMapRenderer.cpp:219-671 */
static void build_line(frame_t *f, int li) {
	static const float pi = M_PI;
	line_t *line = &f->lines[li];
	line_px *pixels = f->pixels + (size_t)li * f->lineRes;
	const int lineResolution = f->lineRes;
	float minPitch = -pi * 0.4999f, maxPitch = pi * 0.4999f;
	for (int k = 2; k < 6; k++)
		clip_plane(f->frustumN[k], line->horizonDir, &minPitch, &maxPitch);

	float minTan = special_tan(minPitch);
	float maxTan = special_tan(maxPitch);
	{
		float minDiff = lineResolution / 10000.f;
		if (maxTan < minTan + minDiff)
			maxTan = minTan + minDiff;
	}
	line->pitchTanMin = minTan;
	line->pitchScale = lineResolution / (maxTan - minTan);
	line->pitchTanMinI = (int)(minTan * 65536.f);
	line->pitchScaleI = (int)(line->pitchScale * 65536.f);

	float dirX = line->horizonDir.x, dirY = line->horizonDir.y;
	if (fabsf(dirY) < 1.e-4f)
		dirY = 1.e-4f;
	if (fabsf(dirX) < 1.e-4f)
		dirX = 1.e-4f;
	float invDirX = 1.f / dirX, invDirY = 1.f / dirY;
	int signX = dirX > 0.f ? 1 : -1, signY = dirY > 0.f ? 1 : -1;
	int invDirXI = abs((int)(invDirX * 256.f)), invDirYI = abs((int)(invDirY * 256.f));
	int dirXI = abs((int)(dirX * 512.f)), dirYI = abs((int)(dirY * 512.f));

	const float cx = f->origin.x, cy = f->origin.y, cz = f->origin.z;
	const int icz = (int)floorf(cz);
	int rx = (int)(cx * 512.f), ry = (int)(cy * 512.f);
	int irx = rx >> 9, iry = ry >> 9; /* int_fast16_t is 64-bit on x86-64: no truncation */

	const float fogDist = 128.f;
	float distance = 1.e-20f;
	float invDist = 1.f / distance;

	xform_t xf;
	xf.transScale = (float)lineResolution / (maxTan - minTan);
	xf.transOffset = -minTan * xf.transScale;
	for (int i = 0; i < lineResolution; i++) {
		pixels[i].combined = 0;
		pixels[i].depth = 10000.f;
	}
	if (minPitch >= maxPitch)
		return;
	for (int i = 0; i < 65; i++)
		xf.zval[i] = (float)i - cz;
	xf.vmax = lineResolution + 0.5f;

	const float zscale = v3dot(line->horizonDir, f->axis[2]);
	const float heightScale = f->axis[2].z;
	float heightScaleVal[65];
	for (int i = 0; i < 65; i++)
		heightScaleVal[i] = (float)i * heightScale;
	const float depthBias = -cz * heightScale;

	int count = 1;
	unsigned cnt2 = (unsigned)(int)(fogDist * 8.f);

	while (distance < fogDist && (--cnt2) > 0) {
		const int oirx = irx, oiry = iry;
		int wallFace;
		unsigned tx, ty;
		/* DDA, :432-522 */
		tx = (unsigned)((signX > 0 ? 512 - (rx & 511) : (rx & 511)) * invDirXI);
		ty = (unsigned)((signY > 0 ? 512 - (ry & 511) : (ry & 511)) * invDirYI);
		if (tx < ty) {
			irx += signX;
			rx = signX > 0 ? irx << 9 : (irx << 9) - 1;
			int dy = (int)(((unsigned)dirYI * tx) >> 17); /* unsigned wrap-around, logical shift */
			ry = signY > 0 ? ry + dy : ry - dy;
			distance += (float)tx * (1.f / 512.f / 256.f);
			wallFace = signX > 0 ? F_NEGX : F_POSX;
		} else {
			iry += signY;
			int dx = (int)(((unsigned)dirXI * ty) >> 17);
			rx = signX > 0 ? rx + dx : rx - dx;
			ry = signY > 0 ? iry << 9 : (iry << 9) - 1;
			distance += (float)ty * (1.f / 512.f / 256.f);
			wallFace = signY > 0 ? F_NEGY : F_POSY;
		}
		const float oldInvDist = invDist;
		invDist = fast_rcp(distance);
		const float medDist = distance * zscale + depthBias;

		/* floor / ceiling of the cell being left, :579-624 */
		{
			uint64_t s = solid_w(f, oirx, oiry);
			for (int z = 0; z < 64; z++) { /* tops ascending: solid with air above (BuildRle :166-172) */
				if (!((s >> z) & 1) || (z > 0 && ((s >> (z - 1)) & 1)))
					continue;
				if (z > icz) {
					int p1 = transform(&xf, invDist, z), p2 = transform(&xf, oldInvDist, z);
					fill_span(pixels, p1, p2, build_line_pixel(f, oirx, oiry, z, F_NEGZ, medDist + heightScaleVal[z]));
				}
			}
			for (int z = 62; z >= 0; z--) { /* bottoms descending: solid with air below; z=63 never (:177-184) */
				if (!((s >> z) & 1) || ((s >> (z + 1)) & 1))
					continue;
				if (z < icz) {
					int p1 = transform(&xf, invDist, z + 1), p2 = transform(&xf, oldInvDist, z + 1);
					fill_span(pixels, p2, p1,
					          build_line_pixel(f, oirx, oiry, z, F_POSZ, medDist + heightScaleVal[z + 1]));
				}
			}
		}
		/* walls of the cell being entered, :627-659 */
		{
			static const int adjx[4] = {1, -1, 0, 0}, adjy[4] = {0, 0, 1, -1};
			uint64_t s = solid_w(f, irx, iry);
			uint64_t a = solid_w(f, (irx & 511) + adjx[wallFace], (iry & 511) + adjy[wallFace]);
			uint64_t walls = s & ~a;
			int savedP = 0, savedZ = 127;
			for (int z = 0; z < 64; z++) {
				if (!((walls >> z) & 1))
					continue;
				int p1 = savedZ == z ? savedP : transform(&xf, invDist, z);
				int p2 = transform(&xf, invDist, z + 1);
				savedZ = z + 1;
				savedP = p2;
				fill_span(pixels, p1, p2, build_line_pixel(f, irx, iry, z, wallFace, medDist + heightScaleVal[z]));
			}
		}
		/* pitch cull, :661-667 */
		if ((--count) == 0) {
			if ((transform(&xf, invDist, 0) >= lineResolution - 1 && icz >= 0) || transform(&xf, invDist, 63) <= 0)
				break;
			count = 4;
		}
	}
}

/* ---- AtanTable / fastATan / fastATan2, SWMapRenderer.cpp:673-742 ---- */
static uint16_t atan_sm[5000], atan_lg[5000], atan_smN[5000], atan_lgN[5000];
static int atan_ready;
static uint16_t atan_to_fixed(float v) {
	v /= (M_PI * 2.f); /* double division, rounded back to float */
	v *= 65536.f;
	int i = (int)v;
	return (uint16_t)(i & 65535);
}
static void atan_init(void) {
	if (atan_ready)
		return;
	for (int i = 0; i < 5000; i++) {
		atan_sm[i] = atan_to_fixed(atanf(i / 4096.f));
		atan_lg[i] = atan_to_fixed(atanf(1.f / ((i + .5f) / 4096.f)));
		atan_smN[i] = atan_to_fixed(-atanf(i / 4096.f));
		atan_lgN[i] = atan_to_fixed(-atanf(1.f / ((i + .5f) / 4096.f)));
	}
	atan_ready = 1;
}
static inline uint16_t fast_atan(float v) {
	if (v < 0.f) {
		if (v > -1.f) {
			v *= -4096.f;
			return atan_smN[(int)v];
		} else {
			v = fast_div(-4096.f, v);
			return atan_lgN[(int)v];
		}
	} else {
		if (v < 1.f) {
			v *= 4096.f;
			return atan_sm[(int)v];
		} else {
			v = fast_div(4096.f, v);
			return atan_lg[(int)v];
		}
	}
}
static inline uint16_t fast_atan2(float y, float x) {
	if (x == 0.f)
		return y > 0.f ? 16384 : (uint16_t)-16384;
	else if (x > 0.f)
		return fast_atan(fast_div(y, x));
	else
		return (uint16_t)(fast_atan(fast_div(y, x)) + 32768);
}

/* SWMapRenderer.cpp:744-1043 for one emulated thread */
static void render_final(const frame_t *f, float yawMin, float yawMax, unsigned numLines, unsigned threadId,
                         unsigned numThreads, int under, unsigned fw, unsigned fh, uint32_t *fb, float *depthBuf) {
	const float fovX = tanf(f->sc.fovX * 0.5f), fovY = tanf(f->sc.fovY * 0.5f);
	const v3 front = f->axis[2], right = f->axis[0], down = f->axis[1];
	const int lineResolution = f->lineRes;
	v3 v1 = v3add(v3sub(front, v3mul(right, fovX)), v3mul(down, fovY));
	const v3 deltaDown = v3mul(V3(-down.x, -down.y, -down.z), fovY * 2.f / (float)fh);
	const v3 deltaRight = v3mul(right, fovX * 2.f / (float)fw * under);
	static const float pi = M_PI;
	const float yawScale = 65536.f / (pi * 2.f);
	const int32_t yawScale2 = (int32_t)(pi * 2.f / (yawMax - yawMin) * 65536.f);
	const int32_t yawMin2 = (int32_t)(yawMin * yawScale);
	const int blockSize = 8, hBlock = 8 / under;
	const v3 deltaDownLarge = v3mul(deltaDown, (float)blockSize);
	const v3 deltaRightLarge = v3mul(deltaRight, (float)hBlock);
	unsigned startX = threadId * fw / numThreads, endX = (threadId + 1) * fw / numThreads;
	startX = (startX / blockSize) * blockSize;
	endX = (endX / blockSize) * blockSize;
	v1 = v3add(v1, v3mul(deltaRight, (float)(startX / under)));

#define CALC_YAW(v) ((int32_t)(((int)fast_atan2((v).y, (v).x) - yawMin2) & 0xffff) << 8)
#define CALC_PITCH(v) ((int32_t)(to_special_tan((v).z * fast_rsqrt((v).x * (v).x + (v).y * (v).y)) * (65536.f * 8192.f)))
#define SHL8SHR8(a) ((int32_t)((uint32_t)(a) << 8) >> 8)

	for (unsigned fx = startX; fx < endX; fx += blockSize) {
		v3 v2 = v1;
		for (unsigned fy = 0; fy < fh; fy += blockSize) {
			uint32_t *fb2 = fb + fx + fy * fw;
			float *db2 = depthBuf + fx + fy * fw;
			if (v2.z > 0.99f || v2.z < -0.99f) {
				/* SlowBlockPath :940-1033 */
				v3 v3_ = v2;
				for (int x = 0; x < blockSize; x += under) {
					v3 v4 = v3_;
					uint32_t *fb3 = fb2 + x;
					float *db3 = db2 + x;
					for (int y = 0; y < blockSize; y++) {
						uint32_t yawIndex = (uint32_t)(((int)fast_atan2(v4.y, v4.x) - yawMin2) & 0xffff);
						yawIndex = (yawIndex * (uint32_t)yawScale2) >> 16;
						yawIndex = (yawIndex * numLines) >> 16;
						const line_t *line = &f->lines[yawIndex];
						float pitch = v4.z * fast_rsqrt(v4.x * v4.x + v4.y * v4.y);
						pitch = to_special_tan(pitch);
						pitch = (pitch - line->pitchTanMin) * line->pitchScale;
						int32_t pitchIndex = (int32_t)pitch;
						pitchIndex &= lineResolution - 1;
						const line_px *pix = &f->pixels[(size_t)yawIndex * lineResolution + pitchIndex];
						for (int k = 0; k < under; k++) {
							fb3[k] = pix->combined;
							db3[k] = pix->depth;
						}
						fb3 += fw;
						db3 += fw;
						v4 = v3add(v4, deltaDown);
					}
					v3_ = v3add(v3_, deltaRight);
				}
			} else {
				/* FastBlockPath :805-937 */
				v3 vr = v3add(v2, deltaRightLarge), vd = v3add(v2, deltaDownLarge);
				v3 vrd = v3add(v3add(v2, deltaRightLarge), deltaDownLarge);
				int32_t yawIndex1 = CALC_YAW(v2), pitch1 = CALC_PITCH(v2);
				int32_t yawIndex2 = CALC_YAW(vr), pitch2 = CALC_PITCH(vr);
				int32_t yawIndex3 = CALC_YAW(vd), pitch3 = CALC_PITCH(vd);
				int32_t yawIndex4 = CALC_YAW(vrd), pitch4 = CALC_PITCH(vrd);
				int32_t yawDiff1 = SHL8SHR8(yawIndex2 - yawIndex1) / hBlock;
				int32_t yawDiff2 = SHL8SHR8(yawIndex4 - yawIndex3) / hBlock;
				int32_t pitchDiff1 = (pitch2 - pitch1) / hBlock;
				int32_t pitchDiff2 = (pitch4 - pitch3) / hBlock;
				int32_t yawIndexA = yawIndex1, yawIndexB = yawIndex3, pitchA = pitch1, pitchB = pitch3;
				for (int x = 0; x < blockSize; x += under) {
					uint32_t *fb3 = fb2 + x;
					float *db3 = db2 + x;
					int32_t yawIndexC = yawIndexA;
					int32_t yawDelta = SHL8SHR8(yawIndexB - yawIndexA) / blockSize;
					int32_t pitchC = pitchA;
					int32_t pitchDelta = (pitchB - pitchA) / blockSize;
					for (int y = 0; y < blockSize; y++) {
						uint32_t yawIndex = (uint32_t)((int32_t)((uint32_t)yawIndexC << 8) >> 16);
						yawIndex = (yawIndex * (uint32_t)yawScale2) >> 16;
						yawIndex = (yawIndex * numLines) >> 16;
						const line_t *line = &f->lines[yawIndex];
						int32_t pitchIndex = pitchC >> 13;
						pitchIndex -= line->pitchTanMinI;
						pitchIndex = (int32_t)(((int64_t)pitchIndex * (int64_t)line->pitchScaleI) >> 32);
						pitchIndex &= lineResolution - 1;
						const line_px *pix = &f->pixels[(size_t)yawIndex * lineResolution + pitchIndex];
						for (int k = 0; k < under; k++) {
							fb3[k] = pix->combined;
							db3[k] = pix->depth;
						}
						fb3 += fw;
						db3 += fw;
						yawIndexC += yawDelta;
						pitchC += pitchDelta;
					}
					yawIndexA += yawDiff1;
					yawIndexB += yawDiff2;
					pitchA += pitchDiff1;
					pitchB += pitchDiff2;
				}
			}
			v2 = v3add(v2, deltaDownLarge);
		}
		v1 = v3add(v1, deltaRightLarge);
	}
}

/* SWUtils.h:68-71 */
static inline int to_fixed8(float v) {
	int i = (int)(v * 255.f + .5f);
	return i > 255 ? 255 : (i < 0 ? 0 : i);
}

/* SWRenderer.cpp:341-516 for one emulated thread; sse2 selects the rounding variant (:422-516) */
static void apply_fog(const orc_scene *sc, const float fogColor[3], float fogDistance, int sse2, unsigned threadId,
                      unsigned numThreads, int fw, int fh, uint32_t *fbAll, float *dbAll) {
	const float fovX = tanf(sc->fovX * 0.5f), fovY = tanf(sc->fovY * 0.5f);
	const float dvx = -fovX * 2.f / (float)(fw / 4);
	const float dvy = -fovY * 2.f / (float)(fh / 4);
	const int fogR = to_fixed8(fogColor[0]), fogG = to_fixed8(fogColor[1]), fogB = to_fixed8(fogColor[2]);
	const uint32_t fog1 = (uint32_t)(fogB + fogR * 0x10000), fog2 = (uint32_t)(fogG * 0x100);
	const float scale = 255.f / fogDistance;
	int startY = (int)((unsigned)fh * threadId / numThreads);
	int endY = (int)((unsigned)fh * (threadId + 1) / numThreads);
	startY &= ~3;
	endY &= ~3;
	float vy = fovY;
	vy += dvy * (startY >> 2);
	uint32_t *fb = fbAll + (size_t)fw * startY;
	float *db = dbAll + (size_t)fw * startY;
	for (int y = startY; y < endY; y += 4) {
		float vx = fovX;
		for (int x = 0; x < fw; x += 4) {
			float depthScale = (1.f + vx * vx + vy * vy);
			depthScale *= fast_rsqrt(depthScale) * scale;
			for (int by = 0; by < 4; by++)
				for (int bx = 0; bx < 4; bx++) {
					uint32_t *p = fb + x + bx + (size_t)by * fw;
					float dist = db[x + bx + (size_t)by * fw] * depthScale;
					int factor;
					if (sse2) {
						/* maxps(dist,0) / minps(.,256) return the second operand on NaN; cvtps2dq rounds to
						   nearest even under the default MXCSR */
						float d = dist > 0.f ? dist : 0.f;
						d = d < 256.f ? d : 256.f;
						factor = (int)lrintf(d);
					} else {
						int t = (int)dist;
						factor = t < 256 ? t : 256;
						factor = 0 < factor ? factor : 0;
					}
					const int factor2 = 256 - factor;
					const uint32_t color = *p;
					if (sse2) {
						/* per-16-bit-lane c*(256-f) + fog*f, >> 8, pack with unsigned saturation; lane 3 = X byte */
						uint32_t b = (((color & 0xff) * factor2 + fogB * factor) & 0xffff) >> 8;
						uint32_t g = ((((color >> 8) & 0xff) * factor2 + fogG * factor) & 0xffff) >> 8;
						uint32_t r = ((((color >> 16) & 0xff) * factor2 + fogR * factor) & 0xffff) >> 8;
						uint32_t a = ((((color >> 24) & 0xff) * factor2) & 0xffff) >> 8;
						*p = b | (g << 8) | (r << 16) | (a << 24);
					} else {
						uint32_t v1 = (color & 0xff00ff) * factor2;
						uint32_t v2 = (color & 0x00ff00) * factor2;
						v1 += fog1 * factor;
						v2 += fog2 * factor;
						v1 &= 0xff00ff00;
						v2 &= 0xff0000;
						*p = (v1 | v2) >> 8;
					}
				}
			vx += dvx;
		}
		vy += dvy;
		fb += (size_t)fw * 4;
		db += (size_t)fw * 4;
	}
}

static __thread line_t *tl_lines;
static __thread line_px *tl_pixels;
static __thread uint32_t *tl_comb;
static __thread float *tl_depth;
static __thread int tl_numLines, tl_lineRes;

int orc_last_lines(int *lineRes, const uint32_t **combined, const float **depth) {
	if (lineRes)
		*lineRes = tl_lineRes;
	size_t n = (size_t)tl_numLines * tl_lineRes;
	tl_comb = realloc(tl_comb, n * 4 + 4);
	tl_depth = realloc(tl_depth, n * 4 + 4);
	for (size_t i = 0; i < n; i++) {
		tl_comb[i] = tl_pixels[i].combined;
		tl_depth[i] = tl_pixels[i].depth;
	}
	if (combined)
		*combined = tl_comb;
	if (depth)
		*depth = tl_depth;
	return tl_numLines;
}

/* EndScene up to the fog (clear + terrain), and the fog pass on its own: orc_models.c draws between the two */
int orc_render_terrain(const uint64_t *solid, const uint32_t *color, const orc_scene *scene, int w, int h, int numerics,
                       int threads, int under, uint32_t *bgrx, float *depth);
void orc_apply_fog(const orc_scene *scene, const float fogColor[3], float fogDist, int numerics, int threads, int w, int h,
                   uint32_t *bgrx, float *depth) {
	if (threads < 1)
		threads = 1;
	if (threads > 32)
		threads = 32;
	for (int t = 0; t < threads; t++)
		apply_fog(scene, fogColor, fogDist, numerics == ORC_NUM_SSE2, t, threads, w, h, bgrx, depth);
}

int orc_render(const uint64_t *solid, const uint32_t *color, const orc_scene *scene, int w, int h,
               const float fogColor[3], float fogDist, int numerics, int threads, int under, uint32_t *bgrx,
               float *depth) {
	int rc = orc_render_terrain(solid, color, scene, w, h, numerics, threads, under, bgrx, depth);
	if (rc)
		return rc;
	orc_apply_fog(scene, fogColor, fogDist, numerics, threads, w, h, bgrx, depth);
	return 0;
}

int orc_render_terrain(const uint64_t *solid, const uint32_t *color, const orc_scene *scene, int w, int h, int numerics,
                       int threads, int under, uint32_t *bgrx, float *depth) {
	static const float pi = M_PI;
	if ((w & 7) || (h & 7) || w <= 0 || h <= 0)
		return -1; /* SWRenderer.cpp:105-107 */
	if (threads < 1)
		threads = 1;
	if (threads > 32)
		threads = 32; /* SWUtils.h:51-53 */
	under = under < 1 ? 1 : (under > 4 ? 4 : under); /* SWMapRenderer.cpp:1110-1111 */
	const int underT = under <= 1 ? 1 : (under <= 2 ? 2 : 4); /* template selection :1160-1169 */
	atan_init();
	g_sse = numerics != ORC_NUM_IEEE;

	frame_t f;
	memset(&f, 0, sizeof(f));
	f.solid = solid;
	f.color = color;
	f.sc = *scene;
	f.numerics = numerics;
	f.origin = V3(scene->origin[0], scene->origin[1], scene->origin[2]);
	for (int i = 0; i < 3; i++)
		f.axis[i] = V3(scene->axis[i][0], scene->axis[i][1], scene->axis[i][2]);

	/* EndScene clear, SWRenderer.cpp:714-717 (depth is NOT cleared) */
	for (size_t i = 0; i < (size_t)w * h; i++)
		bgrx[i] = 0x7f7f7f;

	/* BuildFrustrum, SWRenderer.cpp:207-227 (only the four side-plane normals are used) */
	{
		float xCos = cosf(scene->fovX * .5f), xSin = sinf(scene->fovX * .5f);
		float yCos = cosf(scene->fovY * .5f), ySin = sinf(scene->fovY * .5f);
		f.frustumN[2] = v3sub(v3mul(f.axis[2], xSin), v3mul(f.axis[0], xCos));
		f.frustumN[3] = v3add(v3mul(f.axis[2], xSin), v3mul(f.axis[0], xCos));
		f.frustumN[4] = v3sub(v3mul(f.axis[2], ySin), v3mul(f.axis[1], yCos));
		f.frustumN[5] = v3add(v3mul(f.axis[2], ySin), v3mul(f.axis[1], yCos));
	}

	/* SWMapRenderer::Render :1181-1184: nothing drawn when the eye is inside a solid voxel */
	int px = (int)floorf(f.origin.x), py = (int)floorf(f.origin.y), pz = (int)floorf(f.origin.z);
	int inside = pz >= 64 || (pz >= 0 && ((solid_w(&f, px, py) >> pz) & 1));
	if (!inside) {
		/* RenderInner :1057-1126 */
		float fovX = tanf(scene->fovX * 0.5f), fovY = tanf(scene->fovY * 0.5f);
		float fovDiag = sqrtf(fovX * fovX + fovY * fovY);
		float fovDiagAng = atanf(fovDiag);
		float pitch = asinf(f.axis[2].z);
		float yawMin, yawMax;
		if (fabsf(pitch) >= pi * 0.49f - fovDiagAng) {
			yawMin = 0.f;
			yawMax = pi * 2.f;
		} else {
			float yaw = atan2l(f.axis[2].y, f.axis[2].x);
			yawMin = yaw - pi * .5f;
			yawMax = yaw + pi * .5f;
		}
		float interval = (float)h;
		interval = fovY * 2.f / interval;
		int lineResolution = h;
		for (int i = lineResolution, j = 1; j <= i; j <<= 1)
			lineResolution = j;
		size_t numLines = (size_t)((yawMax - yawMin) / interval);
		numLines /= under;
		if (numLines < 8)
			numLines = 8;
		if (numLines > 65536)
			numLines = 65536;
		f.lineRes = lineResolution;
		f.numLines = (int)numLines;
		tl_lines = f.lines = realloc(tl_lines, sizeof(line_t) * numLines);
		tl_pixels = f.pixels = realloc(tl_pixels, sizeof(line_px) * numLines * lineResolution);
		tl_numLines = (int)numLines;
		tl_lineRes = lineResolution;
		/* line directions :1129-1143 */
		{
			float scl = (yawMax - yawMin) / numLines;
			v3 horiz = V3(cosf(yawMin), sinf(yawMin), 0.f);
			float c = cosf(scl), s = sinf(scl);
			for (size_t i = 0; i < numLines; i++) {
				f.lines[i].horizonDir = horiz;
				float x = horiz.x * c - horiz.y * s;
				float y = horiz.x * s + horiz.y * c;
				horiz.x = x;
				horiz.y = y;
			}
		}
		for (size_t i = 0; i < numLines; i++)
			build_line(&f, (int)i);
		for (int t = 0; t < threads; t++)
			render_final(&f, yawMin, yawMax, (unsigned)numLines, t, threads, underT, w, h, bgrx, depth);
	}
	return 0;
}
