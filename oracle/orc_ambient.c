/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's integer voxel ray walk and of the
 * CPU half of its ray-traced ambient-occlusion volume (see osb_oracle.h for the pinning statement).
 * Plain C, IEEE binary32 (one multiply in double, as the reference's expression promotes), no
 * contraction (built with -ffp-contract=off), same evaluation order as the reference.
 *
 * Paths below are relative to /root/reference/Sources.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "osb_oracle.h"

#define MW 512
#define MH 512
#define MD 64
#define CHUNK 16     /* Draw/GLAmbientShadowRenderer.h:42-43 */
#define NUM_RAYS 16  /* :41 */
#define RAY_LEN 16   /* :44 */
#define PAD 2        /* GLAmbientShadowRenderer.cpp:344 */
#define WIN (CHUNK + 2 * PAD)

static inline uint64_t column(const uint64_t *solid, int x, int y) { return solid[(size_t)(x & (MW - 1)) * MH + (y & (MH - 1))]; }

/* Client/GameMap.h:91-98 IsSolidWrapped */
static inline int solid_wrapped(const uint64_t *solid, int x, int y, int z) {
	if (z < 0)
		return 0;
	if (z >= MD)
		return 1;
	return (int)((column(solid, x, y) >> z) & 1u);
}

/* One axis of the set-up in Client/GameMap.cpp:234-275: step sign, distance to the first cell wall,
 * 1024 x the span on this axis, and the number of cell walls crossed on it. */
static void axis_setup(float p0, float p1, int a, int c, int *d, float *f, float *g, long *cnt) {
	if (c < a) {
		*d = -1;
		*f = p0 - (float)a;
		*g = (p0 - p1) * 1024;
		*cnt += a - c;
	} else if (c != a) {
		*d = 1;
		*f = (float)(a + 1) - p0;
		*g = (p1 - p0) * 1024;
		*cnt += c - a;
	} else {
		*d = 0;
		*f = *g = 0.f;
	}
}

/* Client/GameMap.cpp:208-359 GameMap::CastRay ("faster version" branch, the one that is compiled).
 * Walks at most (long)length cell walls from v0 towards v0 + dir*length; the choice of the next wall is the
 * sign test on three integer cross-product accumulators.  Returns 1 and the block on a hit. */
int orc_castray_one(const uint64_t *solid, const float v0[3], const float dir[3], float length, int32_t hit[3]) {
	float v1[3];
	for (int k = 0; k < 3; k++)
		v1[k] = v0[k] + dir[k] * length;
	int a[3], c[3], d[3];
	float f[3], g[3];
	long cnt = 0;
	for (int k = 0; k < 3; k++) {
		a[k] = (int)floorf(v0[k]);
		c[k] = (int)floorf(v1[k]);
		axis_setup(v0[k], v1[k], a[k], c[k], &d[k], &f[k], &g[k], &cnt);
	}
	int px = (int)floorf(f[0] * g[2] - f[2] * g[0]);
	int py = (int)floorf(f[1] * g[2] - f[2] * g[1]);
	int pz = (int)floorf(f[1] * g[0] - f[0] * g[1]);
	const int ix = (int)floorf(g[0]), iy = (int)floorf(g[1]), iz = (int)floorf(g[2]);
	if (cnt > (long)length)
		cnt = (long)length;

	uint64_t col = column(solid, a[0], a[1]);
	if (a[2] < 0 && d[2] < 0)
		return 0;
	if (a[2] >= MD) { /* :307-309 */
		hit[0] = a[0];
		hit[1] = a[1];
		hit[2] = a[2];
		return 1;
	}
	while (cnt > 0) {
		const int above = a[2] < 0; /* :290-306: above the map nothing is tested and the column is not re-read */
		if (((px | py) >= 0) && a[2] != c[2]) {
			a[2] += d[2];
			px -= ix;
			py -= iy;
			if (!above) {
				if (a[2] < 0 && d[2] < 0)
					return 0;
				if (a[2] >= MD) {
					hit[0] = a[0];
					hit[1] = a[1];
					hit[2] = a[2];
					return 1;
				}
			}
		} else if (pz >= 0 && a[0] != c[0]) {
			a[0] += d[0];
			px += iz;
			pz -= iy;
			if (!above)
				col = column(solid, a[0], a[1]);
		} else {
			a[1] += d[1];
			py += iz;
			pz += ix;
			if (!above)
				col = column(solid, a[0], a[1]);
		}
		if (!above && ((col >> (uint64_t)a[2]) & 1u)) {
			hit[0] = a[0];
			hit[1] = a[1];
			hit[2] = a[2];
			return 1;
		}
		cnt--;
	}
	return 0;
}

void orc_castray(const uint64_t *solid, const float *v0, const float *dir, float length, int32_t *out, size_t n,
                 int threads) {
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
	for (long long i = 0; i < (long long)n; i++) {
		int32_t hb[3] = {0, 0, 0};
		const int h = orc_castray_one(solid, v0 + i * 3, dir + i * 3, length, hb);
		out[i * 4] = h;
		out[i * 4 + 1] = h ? hb[0] : 0;
		out[i * 4 + 2] = h ? hb[1] : 0;
		out[i * 4 + 3] = h ? hb[2] : 0;
	}
}

/* Draw/GLAmbientShadowRenderer.cpp:130-169 Evaluate: 16 rays of length 16 from the voxel centre, ray i is
 * rays[i] mirrored by the bits of i & 7; a hit contributes |hit centre - origin|^2 / 15^2 (clamped to 1), a miss 1. */
float orc_ao_evaluate(const uint64_t *solid, const float rays[NUM_RAYS * 3], int x, int y, int z) {
	float sum = 0.0f;
	float pos[3] = {(float)x, (float)y, (float)z};
	for (int k = 0; k < 3; k++)
		pos[k] += 0.5f;
	for (int i = 0; i < NUM_RAYS; i++) {
		float dir[3] = {rays[i * 3], rays[i * 3 + 1], rays[i * 3 + 2]};
		for (int k = 0; k < 3; k++)
			if ((i & 7) & (1 << k))
				dir[k] = -dir[k];
		int32_t hb[3];
		float brightness = 1.f;
		if (orc_castray_one(solid, pos, dir, (float)RAY_LEN, hb)) {
			const float dx = ((float)hb[0] + .5f) - pos[0], dy = ((float)hb[1] + .5f) - pos[1], dz = ((float)hb[2] + .5f) - pos[2];
			const float dist = dx * dx + dy * dy + dz * dz;
			/* `dist * (1.0 / float(15 * 15))`: the divisor expression is double, so is the product */
			brightness = (float)((double)dist * (1.0 / (double)(float)((RAY_LEN - 1) * (RAY_LEN - 1))));
			if (brightness > 1.f)
				brightness = 1.f;
		}
		sum += brightness;
	}
	const float s = sum * (2.f / (float)NUM_RAYS);
	return s < 1.0f ? s : 1.0f; /* std::min(s, 1.0f) */
}

void orc_ao_evaluate_box(const uint64_t *solid, const float rays[NUM_RAYS * 3], int x0, int y0, int z0, int x1, int y1,
                         int z1, float *out, int threads) {
	const int nx = x1 - x0 + 1, ny = y1 - y0 + 1, nz = z1 - z0 + 1;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic, 4)
	for (int row = 0; row < nz * ny; row++) {
		const int z = z0 + row / ny, y = y0 + row % ny;
		for (int x = x0; x <= x1; x++)
			out[(size_t)row * nx + (x - x0)] = orc_ao_evaluate(solid, rays, x, y, z);
	}
}

/* chunk state: 4096 x {dirty, minX, minY, minZ, maxX, maxY, maxZ}, index (cx + cy*32)*4 + cz
 * (GLAmbientShadowRenderer.h:55-66: a new chunk is dirty over its whole 16^3) */
void orc_ao_state_init(int32_t *state) {
	for (int i = 0; i < 4096; i++) {
		int32_t *s = state + i * 7;
		s[0] = 1;
		s[1] = s[2] = s[3] = 0;
		s[4] = s[5] = s[6] = CHUNK - 1;
	}
}

/* Draw/GLAmbientShadowRenderer.cpp:171-237 GameMapChanged + Invalidate: the +-16 box around the edit, z clamped to
 * the map, spread over the (x,y-wrapped) chunks it touches; first touch assigns the box, later ones grow it. */
void orc_ao_invalidate(int32_t *state, int x, int y, int z) {
	int lo[3] = {x - RAY_LEN, y - RAY_LEN, z - RAY_LEN}, hi[3] = {x + RAY_LEN, y + RAY_LEN, z + RAY_LEN};
	if (lo[2] < 0)
		lo[2] = 0;
	if (hi[2] > MD - 1)
		hi[2] = MD - 1;
	if (lo[0] > hi[0] || lo[1] > hi[1] || lo[2] > hi[2])
		return;
	for (int cx = lo[0] >> 4; cx <= hi[0] >> 4; cx++)
		for (int cy = lo[1] >> 4; cy <= hi[1] >> 4; cy++)
			for (int cz = lo[2] >> 4; cz <= hi[2] >> 4; cz++) {
				int32_t *s = state + (((cx & 31) + (cy & 31) * 32) * 4 + cz) * 7;
				const int org[3] = {cx * CHUNK, cy * CHUNK, cz * CHUNK};
				int bmin[3], bmax[3];
				for (int k = 0; k < 3; k++) {
					bmin[k] = lo[k] - org[k] > 0 ? lo[k] - org[k] : 0;
					bmax[k] = hi[k] - org[k] < CHUNK - 1 ? hi[k] - org[k] : CHUNK - 1;
				}
				if (!s[0]) {
					for (int k = 0; k < 3; k++) {
						s[1 + k] = bmin[k];
						s[4 + k] = bmax[k];
					}
					s[0] = 1;
				} else {
					for (int k = 0; k < 3; k++) {
						if (bmin[k] < s[1 + k])
							s[1 + k] = bmin[k];
						if (bmax[k] > s[4 + k])
							s[4 + k] = bmax[k];
					}
				}
			}
}

/* bit 0 = solid, bit 1 = "contact": some neighbour of the 3x3x3 shell is solid.  The reference's list
 * (GLAmbientShadowRenderer.cpp:383-409) names (x-1,y-1,z-1) twice and never (x,y-1,z-1); kept as is. */
static uint8_t ao_flags(const uint64_t *solid, int x, int y, int z) {
	int contact = 0;
	for (int dx = -1; dx <= 1; dx++)
		for (int dy = -1; dy <= 1; dy++)
			for (int dz = -1; dz <= 1; dz++) {
				if ((dx == 0 && dy == 0 && dz == 0) || (dx == 0 && dy == -1 && dz == -1))
					continue;
				contact |= solid_wrapped(solid, x + dx, y + dy, z + dz);
			}
	return (uint8_t)(solid_wrapped(solid, x, y, z) | (contact << 1));
}

/* shouldBlur (:457-459): the neighbour is air, and it is a contact voxel or this one is not */
static inline int should_blur(uint8_t self, uint8_t nb) { return ((nb & 1) | ((~self | nb) & 2)) == 2; }

/* Draw/GLAmbientShadowRenderer.cpp:334-507 UpdateChunk.  data = this chunk's float[16][16][16][2] (AO term, air flag).
 * Window = dirty box grown by 2; AO per air voxel, under-shadowing compensation d *= d*d + 1 - d, then two rounds of
 * in-place 3-tap blurs along x, y, z (each voxel sees its already-blurred predecessor), solid voxels untouched. */
void orc_ao_update_chunk(const uint64_t *solid, const float rays[NUM_RAYS * 3], int32_t *state, int cx, int cy, int cz,
                         float *data) {
	int32_t *s = state + ((cx + cy * 32) * 4 + cz) * 7;
	if (!s[0])
		return;
	static const float divider[3] = {1.0f, 1.0f / 2.0f, 1.0f / 3.0f};
	static __thread float w[WIN][WIN][WIN];
	static __thread uint8_t wf[WIN][WIN][WIN];
	const int ox = cx * CHUNK - PAD, oy = cy * CHUNK - PAD, oz = cz * CHUNK - PAD;
	const int x0 = s[1], y0 = s[2], z0 = s[3], x1 = s[4] + 2 * PAD, y1 = s[5] + 2 * PAD, z1 = s[6] + 2 * PAD;
	for (int z = z0; z <= z1; z++)
		for (int y = y0; y <= y1; y++)
			for (int x = x0; x <= x1; x++) {
				const int wx = x + ox, wy = y + oy, wz = z + oz;
				wf[z][y][x] = ao_flags(solid, wx, wy, wz);
				float d = (wf[z][y][x] & 1) ? 0.0f : orc_ao_evaluate(solid, rays, wx, wy, wz);
				d *= d * d + 1.0f - d; /* :417-423 */
				w[z][y][x] = d;
			}
	for (int pass = 0; pass < 2; pass++) {
		for (int z = z0; z <= z1; z++)
			for (int y = y0; y <= y1; y++)
				for (int x = x0 + 1; x < x1; x++) {
					if (wf[z][y][x] & 1)
						continue;
					const int m1 = should_blur(wf[z][y][x], wf[z][y][x - 1]), m2 = should_blur(wf[z][y][x], wf[z][y][x + 1]);
					w[z][y][x] = (w[z][y][x] + (m1 ? w[z][y][x - 1] : 0.0f) + (m2 ? w[z][y][x + 1] : 0.0f)) * divider[m1 + m2];
				}
		for (int z = z0; z <= z1; z++)
			for (int y = y0 + 1; y < y1; y++)
				for (int x = x0; x <= x1; x++) {
					if (wf[z][y][x] & 1)
						continue;
					const int m1 = should_blur(wf[z][y][x], wf[z][y - 1][x]), m2 = should_blur(wf[z][y][x], wf[z][y + 1][x]);
					w[z][y][x] = (w[z][y][x] + (m1 ? w[z][y - 1][x] : 0.0f) + (m2 ? w[z][y + 1][x] : 0.0f)) * divider[m1 + m2];
				}
		for (int z = z0 + 1; z < z1; z++)
			for (int y = y0; y <= y1; y++)
				for (int x = x0; x <= x1; x++) {
					if (wf[z][y][x] & 1)
						continue;
					const int m1 = should_blur(wf[z][y][x], wf[z - 1][y][x]), m2 = should_blur(wf[z][y][x], wf[z + 1][y][x]);
					w[z][y][x] = (w[z][y][x] + (m1 ? w[z - 1][y][x] : 0.0f) + (m2 ? w[z + 1][y][x] : 0.0f)) * divider[m1 + m2];
				}
	}
	for (int z = s[3]; z <= s[6]; z++)
		for (int y = s[2]; y <= s[5]; y++)
			for (int x = s[1]; x <= s[4]; x++) {
				float *o = data + (((size_t)z * CHUNK + y) * CHUNK + x) * 2;
				o[0] = w[z + PAD][y + PAD][x + PAD];
				o[1] = (wf[z + PAD][y + PAD][x + PAD] & 1) ? 0.0f : 1.0f;
			}
	s[0] = 0;
}
