/*
 * TEST INFRASTRUCTURE ONLY (see osb_oracle.h) -- CPU restatement of the reference's map code:
 * VXL decode, CastRay2, SW renderer RLE records.  Compile with -ffp-contract=off, no -march flags
 * (the reference is plain x86-64 code without FMA).
 */
#include "osb_oracle.h"

#include <math.h>
#include <string.h>

#define W 512
#define H 512
#define D 64

static inline uint64_t solid_wrapped(const uint64_t *solid, int x, int y) {
	return solid[(size_t)(x & (W - 1)) * H + (y & (H - 1))]; /* GameMap.h:87-89 */
}
static inline int is_solid_wrapped(const uint64_t *solid, int x, int y, int z) {
	/* GameMap.h:91-98 */
	if (z < 0)
		return 0;
	if (z >= D)
		return 1;
	return (int)((solid_wrapped(solid, x, y) >> z) & 1ull);
}

/* GameMap.cpp:509-517 swapColor: BGRA file order -> 0x64BBGGRR */
static uint32_t swap_color(uint32_t col) {
	uint32_t b0 = col & 0xff, b2 = (col >> 16) & 0xff;
	col = (col & 0xff00ff00u) | (b0 << 16) | b2;
	return (col & 0xffffff) | (100u * 0x1000000u);
}

/* GameMap.cpp:519-596 */
int orc_vxl_load(const uint8_t *data, size_t len, uint64_t *solid, uint32_t *color) {
	size_t pos = 0;
	memset(color, 0, sizeof(uint32_t) * W * H * D);
#define NEED(p)                                                                                    \
	do {                                                                                           \
		if ((p) >= len)                                                                            \
			return -1;                                                                             \
	} while (0)
#define RD32(p) ((uint32_t)data[p] | ((uint32_t)data[(p) + 1] << 8) | ((uint32_t)data[(p) + 2] << 16) | ((uint32_t)data[(p) + 3] << 24))
	for (int y = 0; y < H; y++)
		for (int x = 0; x < W; x++) {
			uint64_t s = ~0ull;
			uint32_t *c = color + ((size_t)x * H + y) * D;
			int z = 0;
			for (;;) {
				NEED(pos + 3);
				int n4 = (int8_t)data[pos];
				int top_start = (int8_t)data[pos + 1];
				int top_end = (int8_t)data[pos + 2];
				for (int i = z; i < top_start; i++)
					s &= ~(1ull << i);
				size_t co = pos + 4;
				for (z = top_start; z <= top_end; z++) {
					NEED(co + 3);
					c[z] = swap_color(RD32(co));
					s |= 1ull << z;
					co += 4;
				}
				if (top_end == 62) {
					c[63] = c[62];
					s |= 1ull << 63;
				}
				int len_bottom = top_end - top_start + 1;
				if (n4 == 0) {
					pos += 4 * (size_t)(len_bottom + 1);
					break;
				}
				int len_top = (n4 - 1) - len_bottom;
				pos += (size_t)n4 * 4;
				NEED(pos + 3);
				int bottom_end = (int8_t)data[pos + 3];
				int bottom_start = bottom_end - len_top;
				for (z = bottom_start; z < bottom_end; z++) {
					NEED(co + 3);
					c[z] = swap_color(RD32(co));
					s |= 1ull << z;
					co += 4;
				}
				if (bottom_end == 63) {
					c[63] = c[62];
					s |= 1ull << 63;
				}
			}
			solid[(size_t)x * H + y] = s;
		}
	return 0;
#undef NEED
#undef RD32
}

/* GameMap.cpp:65-83 (no wrap: the map edge has no neighbour) */
static int is_solid(const uint64_t *solid, int x, int y, int z) { return (int)((solid[(size_t)x * H + y] >> z) & 1); }
static int is_surface(const uint64_t *solid, int x, int y, int z) {
	if (!is_solid(solid, x, y, z))
		return 0;
	if (z == 0)
		return 1;
	if (x > 0 && !is_solid(solid, x - 1, y, z))
		return 1;
	if (x < W - 1 && !is_solid(solid, x + 1, y, z))
		return 1;
	if (y > 0 && !is_solid(solid, x, y - 1, z))
		return 1;
	if (y < H - 1 && !is_solid(solid, x, y + 1, z))
		return 1;
	if (!is_solid(solid, x, y, z - 1))
		return 1;
	if (z < D - 1 && !is_solid(solid, x, y, z + 1))
		return 1;
	return 0;
}

/* GameMap.cpp:93-160 GameMap::Save.  Returns the stream length; writes at most cap bytes. */
long long orc_vxl_save(const uint64_t *solid, const uint32_t *color, uint8_t *out, size_t cap) {
	size_t n = 0;
#define PUT(b)                                                                                     \
	do {                                                                                           \
		if (n < cap)                                                                               \
			out[n] = (uint8_t)(b);                                                                 \
		n++;                                                                                       \
	} while (0)
	for (int y = 0; y < H; y++)
		for (int x = 0; x < W; x++) {
			const uint32_t *c = color + ((size_t)x * H + y) * D;
			int k = 0;
			while (k < D) {
				int z;
				int air_start = k;
				while (k < D && !is_solid(solid, x, y, k))
					++k;
				int top_colors_start = k;
				while (k < D && is_surface(solid, x, y, k))
					++k;
				int top_colors_end = k;
				while (k < D && is_solid(solid, x, y, k) && !is_surface(solid, x, y, k))
					++k;
				int bottom_colors_start = k;
				z = k;
				while (z < D && is_surface(solid, x, y, z))
					++z;
				if (z != D) {
					while (is_surface(solid, x, y, k))
						++k;
				}
				int bottom_colors_end = k;
				int top_colors_len = top_colors_end - top_colors_start;
				int bottom_colors_len = bottom_colors_end - bottom_colors_start;
				int colors = top_colors_len + bottom_colors_len;
				if (k == D)
					PUT(0);
				else
					PUT(colors + 1);
				PUT(top_colors_start);
				PUT(top_colors_end - 1);
				PUT(air_start);
				for (z = 0; z < top_colors_len; ++z) {
					uint32_t col = c[top_colors_start + z];
					PUT(col >> 16);
					PUT(col >> 8);
					PUT(col >> 0);
					PUT(col >> 24);
				}
				for (z = 0; z < bottom_colors_len; ++z) {
					uint32_t col = c[bottom_colors_start + z];
					PUT(col >> 16);
					PUT(col >> 8);
					PUT(col >> 0);
					PUT(col >> 24);
				}
			}
		}
#undef PUT
	return (long long)n;
}

/* GameMap.cpp:361-507, one ray */
static void castray2_one(const uint64_t *solid, const float *v0, const float *d, int maxSteps,
                         orc_rayhit *r) {
	memset(r, 0, sizeof(*r));
	float dx = d[0], dy = d[1], dz = d[2];
	/* Core/Math.h:369-374 Normalize */
	float scale = sqrtf(dx * dx + dy * dy + dz * dz);
	if (scale != 0.f)
		scale = 1.f / scale;
	dx *= scale;
	dy *= scale;
	dz *= scale;
	int ix = (int)floorf(v0[0]), iy = (int)floorf(v0[1]), iz = (int)floorf(v0[2]);
	if (is_solid_wrapped(solid, ix, iy, iz)) { /* :377-384 */
		r->hit = 1;
		r->startSolid = 1;
		r->hitPos[0] = v0[0];
		r->hitPos[1] = v0[1];
		r->hitPos[2] = v0[2];
		r->hitBlock[0] = ix;
		r->hitBlock[1] = iy;
		r->hitBlock[2] = iz;
		return;
	}
	float fx = dx > 0.f ? (float)(ix + 1) - v0[0] : v0[0] - (float)ix; /* :386-400 */
	float fy = dy > 0.f ? (float)(iy + 1) - v0[1] : v0[1] - (float)iy;
	float fz = dz > 0.f ? (float)(iz + 1) - v0[2] : v0[2] - (float)iz;
	float invX = dx, invY = dy, invZ = dz; /* :402-411 */
	if (invX != 0.f)
		invX = 1.f / fabsf(invX);
	if (invY != 0.f)
		invY = 1.f / fabsf(invY);
	if (invZ != 0.f)
		invZ = 1.f / fabsf(invZ);
	for (int i = 0; i < maxSteps; i++) { /* :413-501 */
		int nx = ix, ny = iy, nz = iz, has = 0;
		float t = 0.f;
		if (invX != 0.f) {
			nx = ix + (dx > 0.f ? 1 : -1);
			t = fx * invX;
			has = 1;
		}
		if (invY != 0.f) {
			float tt = fy * invY;
			if (!has || tt < t) {
				nx = ix;
				ny = iy + (dy > 0.f ? 1 : -1);
				t = tt;
				has = 2;
			}
		}
		if (invZ != 0.f) {
			float tt = fz * invZ;
			if (!has || tt < t) {
				nx = ix;
				ny = iy;
				nz = iz + (dz > 0.f ? 1 : -1);
				t = tt;
				has = 3;
			}
		}
		if (has == 1)
			fx = 1.f;
		else
			fx -= fabsf(dx) * t;
		if (has == 2)
			fy = 1.f;
		else
			fy -= fabsf(dy) * t;
		if (has == 3)
			fz = 1.f;
		else
			fz -= fabsf(dz) * t;
		r->hitBlock[0] = nx;
		r->hitBlock[1] = ny;
		r->hitBlock[2] = nz;
		r->normal[0] = ix - nx;
		r->normal[1] = iy - ny;
		r->normal[2] = iz - nz;
		if (is_solid_wrapped(solid, nx, ny, nz)) {
			r->hitPos[0] = dx > 0.f ? (float)(nx + 1) - fx : (float)nx + fx;
			r->hitPos[1] = dy > 0.f ? (float)(ny + 1) - fy : (float)ny + fy;
			r->hitPos[2] = dz > 0.f ? (float)(nz + 1) - fz : (float)nz + fz;
			r->hit = 1;
			return;
		}
		ix = nx;
		iy = ny;
		iz = nz;
	}
	r->hitPos[0] = v0[0]; /* :503-506 */
	r->hitPos[1] = v0[1];
	r->hitPos[2] = v0[2];
}

void orc_castray2(const uint64_t *solid, const float *v0, const float *dir, int maxSteps, orc_rayhit *out,
                  size_t n, int threads) {
	if (threads < 1)
		threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static)
	for (long long i = 0; i < (long long)n; i++)
		castray2_one(solid, v0 + i * 3, dir + i * 3, maxSteps, out + i);
}

/* SWMapRenderer.cpp:142-203 */
int orc_build_rle(const uint64_t *solid, int x, int y, int8_t *out) {
	int n = 10; /* five uint16 header slots */
	memset(out, 0, 10);
	uint64_t smap = solid_wrapped(solid, x, y);
	uint64_t adjs[4] = {solid_wrapped(solid, x + 1, y), solid_wrapped(solid, x - 1, y),
	                    solid_wrapped(solid, x, y + 1), solid_wrapped(solid, x, y - 1)};
	int old = 0;
	for (int z = 0; z < 64; z++) {
		int b = (int)((smap >> z) & 1);
		if (b && !old)
			out[n++] = (int8_t)z;
		old = b;
	}
	out[n++] = -1;
	uint16_t hdr = (uint16_t)n;
	memcpy(out + 0, &hdr, 2);
	old = 1;
	for (int z = 63; z >= 0; z--) {
		int b = (int)((smap >> z) & 1);
		if (b && !old)
			out[n++] = (int8_t)z;
		old = b;
	}
	out[n++] = -1;
	for (int k = 0; k < 4; k++) {
		hdr = (uint16_t)n;
		memcpy(out + 2 * (k + 1), &hdr, 2);
		for (int z = 0; z < 64; z++)
			if (((smap >> z) & 1) && !((adjs[k] >> z) & 1))
				out[n++] = (int8_t)z;
		out[n++] = -1;
	}
	while (n & 3)
		out[n++] = 42;
	return n;
}
