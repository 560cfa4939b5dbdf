/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's sun-shadow bitmap and radiosity
 * evaluation (see osb_oracle.h for the pinning statement).  Plain C, IEEE binary32, no contraction
 * (built with -ffp-contract=off), same evaluation order as the reference so results are bit-identical.
 *
 * Paths below are relative to /root/reference/Sources.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

#include "osb_oracle.h"

#define MW 512
#define MH 512
#define MD 64
#define ENVELOPE 6 /* Draw/GLRadiosityRenderer.h:44 */

static inline int solid_at(const uint64_t *solid, int x, int y, int z) { return (int)((solid[(size_t)x * MH + y] >> z) & 1u); }
static inline uint32_t color_at(const uint32_t *color, int x, int y, int z) { return color[((size_t)x * MH + y) * MD + z]; }

/* Draw/GLMapShadowRenderer.cpp:177-197 BuildPixel: 6-bit colour, depth in the top byte, bit 7 = side face */
static uint32_t shadow_pack(int distance, uint32_t col, int side) {
	const uint32_t r = (col & 0xff) >> 2, g = ((col >> 8) & 0xff) >> 2, b = ((col >> 16) & 0xff) >> 2;
	return r + (g << 8) + (b << 16) + ((uint32_t)distance << 24) + ((uint32_t)(side ? 1 : 0) << 7);
}

/* Draw/GLMapShadowRenderer.cpp:199-218 GeneratePixel: walk the (0,+1,+1) sun diagonal from (x,y,0);
 * at each z test the voxel itself (top face, "z-plane hit") and then the voxel one step in +y
 * ("y-plane hit", recorded at depth z+1); the water floor z = 63 never occludes. */
static uint32_t shadow_pixel(const uint64_t *solid, const uint32_t *color, int x, int y) {
	for (int z = 0; z < MD; z++) {
		if (z < 63 && solid_at(solid, x, y, z))
			return shadow_pack(z, color_at(color, x, y, z), 0);
		y = (y + 1 == MH) ? 0 : y + 1;
		if (z < 63 && solid_at(solid, x, y, z))
			return shadow_pack(z + 1, color_at(color, x, y, z), 1);
	}
	return shadow_pack(64, color_at(color, x, y, 63), 0);
}

void orc_shadow_bitmap(const uint64_t *solid, const uint32_t *color, uint32_t *out) {
#pragma omp parallel for schedule(static)
	for (int y = 0; y < MH; y++)
		for (int x = 0; x < MW; x++)
			out[(size_t)y * MW + x] = shadow_pixel(solid, color, x, y);
}

uint32_t orc_shadow_pixel(const uint64_t *solid, const uint32_t *color, int x, int y) {
	return shadow_pixel(solid, color, x & (MW - 1), y & (MH - 1));
}

/* Draw/GLRadiosityRenderer.cpp:162-275 Evaluate.  out = base.rgb, x.rgb, y.rgb, z.rgb */
void orc_radiosity_evaluate(const uint32_t *bitmap, int vx, int vy, int vz, float out[12]) {
	float acc[12];
	for (int i = 0; i < 12; i++)
		acc[i] = 0.f;
	const float px = (float)vx + .5f, py = (float)vy + .5f, pz = (float)vz + .5f;
	const int cx = vx, cy = vy - vz; /* voxel centre in bitmap space */

	for (int dx = -ENVELOPE; dx <= ENVELOPE; dx++) {
		const uint32_t *column = bitmap + ((cx + dx) & (MW - 1));
		for (int dy = -ENVELOPE; dy <= ENVELOPE; dy++) {
			const uint32_t pixel = column[(size_t)MW * ((cy + dy) & (MH - 1))];
			const int depth = (int)(pixel >> 24);
			/* world position of the lit face this shadow-map texel stands for (not wrapped) */
			const int wx = cx + dx, wy = cy + dy + depth, wz = depth;
			float fx, fy, fz, facing;
			if (pixel & 0x80u) { /* -y facing side plane */
				if (wy <= vy)
					continue;
				fx = px - ((float)wx + .5f);
				fy = py - (float)wy;
				fz = pz - ((float)wz - .5f);
				facing = -fy;
			} else { /* -z facing top plane */
				if (wz <= vz)
					continue;
				fx = px - ((float)wx + .5f);
				fy = py - ((float)wy + .5f);
				fz = pz - (float)wz;
				facing = -fz;
			}
			/* Vector3::GetLength: sqrtf(x*x + y*y + z*z), Core/Math.h */
			const float len = sqrtf(fx * fx + fy * fy + fz * fz);
			const float invLen = 1.f / len;
			const float invSmooth = 1.f / (len + .4f);
			float intensity = facing * invLen;
			intensity *= invSmooth;
			intensity *= invSmooth;
			const float s = -invLen; /* normDiff = diff * -invDiffLen */
			const float nx = fx * s, ny = fy * s, nz = fz * s;
			float c[3];
			c[0] = (float)(pixel & 0x3f);
			c[1] = (float)((pixel >> 8) & 0x3f);
			c[2] = (float)((pixel >> 16) & 0x3f);
			for (int k = 0; k < 3; k++) {
				c[k] *= intensity;
				acc[k] += c[k];
				acc[3 + k] += c[k] * nx;
				acc[6 + k] += c[k] * ny;
				acc[9 + k] += c[k] * nz;
			}
		}
	}
	const float scale = 0.1f / 64.f;
	for (int i = 0; i < 12; i++)
		out[i] = acc[i] * scale;
}

/* Draw/GLRadiosityRenderer.cpp:456-463 CompressDynamicRange */
static float compress_range(float v, int mode) {
	if (mode >= 2)
		return v;
	return v >= 0.f ? sqrtf(v) : -sqrtf(-v);
}

/* one 10-bit channel of EncodeValue, Draw/GLRadiosityRenderer.cpp:465-515 */
static uint32_t encode_channel(float c) {
	float v = c * 1023.f + .5f;
	if (v > 1023.2f)
		v = 1023.2f;
	if (v < 0.f)
		v = 0.f;
	int iv = (int)(unsigned int)v; /* NaN never reaches here: both compares above are false for NaN and
	                                  (unsigned)NaN is what the reference would also produce on x86 */
	if (iv > 1023)
		iv = 1023;
	if (iv < 0)
		iv = 0;
	return (uint32_t)iv;
}

uint32_t orc_radiosity_encode(const float v[3], int mode) {
	float c[3];
	for (int k = 0; k < 3; k++) {
		c[k] = compress_range(v[k], mode);
		c[k] *= .5f;
		c[k] += .5f;
		c[k] *= 1022.f / 1023.f;
	}
	return 0xC0000000u | (encode_channel(c[0]) << 20) | (encode_channel(c[1]) << 10) | encode_channel(c[2]);
}

/* inner loop of UpdateChunk, Draw/GLRadiosityRenderer.cpp:517-543, over an arbitrary box */
void orc_radiosity_box(const uint32_t *bitmap, int x0, int y0, int z0, int x1, int y1, int z1, int mode,
                       float *f32, uint32_t *enc, int threads) {
	const int nx = x1 - x0 + 1, ny = y1 - y0 + 1, nz = z1 - z0 + 1;
	if (nx <= 0 || ny <= 0 || nz <= 0)
		return;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic, 4)
	for (int row = 0; row < nz * ny; row++) {
		const int z = z0 + row / ny, y = y0 + row % ny;
		for (int x = x0; x <= x1; x++) {
			float r[12];
			orc_radiosity_evaluate(bitmap, x, y, z, r);
			const size_t o = (size_t)row * nx + (size_t)(x - x0);
			if (f32)
				for (int i = 0; i < 12; i++)
					f32[o * 12 + i] = r[i];
			if (enc)
				for (int i = 0; i < 4; i++)
					enc[o * 4 + i] = orc_radiosity_encode(r + i * 3, mode);
		}
	}
}

/* ---- invalidation (chunk dirty boxes) -------------------------------------------------------
 * Draw/GLRadiosityRenderer.cpp:277-336 GameMapChanged/Invalidate.  state: 4096 chunks x 7 ints
 * {dirty, minX, minY, minZ, maxX, maxY, maxZ}, chunk index (cx + cy*32)*4 + cz (GLRadiosityRenderer.h:64-72). */
void orc_radiosity_invalidate(int32_t *state, int x, int y, int z) {
	int minX = x - ENVELOPE, minY = y - ENVELOPE, minZ = z - ENVELOPE;
	int maxX = x + ENVELOPE, maxY = y + ENVELOPE, maxZ = z + ENVELOPE;
	if (minZ < 0)
		minZ = 0;
	if (maxZ > MD - 1)
		maxZ = MD - 1;
	if (minX > maxX || minY > maxY || minZ > maxZ)
		return;
	for (int cx = minX >> 4; cx <= (maxX >> 4); cx++)
		for (int cy = minY >> 4; cy <= (maxY >> 4); cy++)
			for (int cz = minZ >> 4; cz <= (maxZ >> 4); cz++) {
				int32_t *c = state + (size_t)(((cx & 31) + (cy & 31) * 32) * 4 + cz) * 7;
				int lo[3] = {minX - cx * 16, minY - cy * 16, minZ - cz * 16};
				int hi[3] = {maxX - cx * 16, maxY - cy * 16, maxZ - cz * 16};
				for (int k = 0; k < 3; k++) {
					if (lo[k] < 0)
						lo[k] = 0;
					if (hi[k] > 15)
						hi[k] = 15;
				}
				if (!c[0]) {
					for (int k = 0; k < 3; k++) {
						c[1 + k] = lo[k];
						c[4 + k] = hi[k];
					}
					c[0] = 1;
				} else {
					for (int k = 0; k < 3; k++) {
						if (lo[k] < c[1 + k])
							c[1 + k] = lo[k];
						if (hi[k] > c[4 + k])
							c[4 + k] = hi[k];
					}
				}
			}
}

/* Draw/GLMapShadowRenderer.cpp:78-108,220-229: an edit at (x,y,z) marks shadow pixels (x, y-z) and
 * (x, y-z-1); Update() regenerates every pixel of the 32-pixel words holding marked pixels and, for each
 * pixel whose value changed, invalidates radiosity around the new and the old lit position.
 * words: 512*16 dirty words (bit = x&31 of word (x>>5) + y*16), consumed (cleared) here.
 * Returns the number of changed pixels. */
int orc_shadow_update(const uint64_t *solid, const uint32_t *color, uint32_t *bitmap, uint32_t *words,
                      int32_t *radState) {
	int changed = 0;
	for (int i = 0; i < MH * (MW / 32); i++) {
		if (!words[i])
			continue;
		const int y = i / (MW / 32), x0 = (i - y * (MW / 32)) * 32;
		for (int j = 0; j < 32; j++) {
			const uint32_t nv = shadow_pixel(solid, color, x0 + j, y);
			uint32_t *cur = bitmap + (size_t)i * 32 + j;
			if (*cur != nv) {
				if (radState) {
					int dist = (int)(nv >> 24);
					orc_radiosity_invalidate(radState, x0 + j, (y + dist) & (MH - 1), dist);
					dist = (int)(*cur >> 24);
					orc_radiosity_invalidate(radState, x0 + j, (y + dist) & (MH - 1), dist);
				}
				*cur = nv;
				changed++;
			}
		}
		words[i] = 0;
	}
	return changed;
}

void orc_shadow_mark(uint32_t *words, int x, int y, int z) {
	for (int k = 0; k < 2; k++) {
		const int px = x & (MW - 1), py = (y - z - k) & (MH - 1);
		words[(px >> 5) + py * (MW / 32)] |= 1u << (px & 31);
	}
}
