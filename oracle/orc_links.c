/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the floating-block detection and of the minimap pixel.
 * Pinned: tests/test_oracle_links.py compares it with the unmodified reference compiled in place (oracle/_ref).
 *
 * Follows, in /root/reference/Sources:
 *   Client/GameMapWrapper.cpp:35-46    constructor (link map, one byte per voxel)
 *   Client/GameMapWrapper.cpp:50-102   Rebuild      -> orc_links_rebuild
 *   Client/GameMapWrapper.cpp:104-193  AddBlock     -> orc_links_add_block
 *   Client/GameMapWrapper.cpp:199-361  RemoveBlocks -> orc_links_remove_blocks
 *   Draw/SWFlatMapRenderer.cpp:94-105  GeneratePixel -> orc_minimap
 */
#include "osb_oracle.h"

#include <stdlib.h>
#include <string.h>

enum { W = 512, H = 512, D = 64 };
enum { L_INVALID = 0, L_ROOT, L_NEGX, L_POSX, L_NEGY, L_POSY, L_NEGZ, L_POSZ, L_MARKED };

typedef struct {
	short x, y, z;
} cell_t;

struct orc_links {
	uint8_t *link; /* [(x * H + y) * D + z] */
	cell_t *queue; /* FIFO, every cell enters at most a bounded number of times per call: ring buffer */
	size_t qcap, qhead, qtail;
};

static inline int solid_at(const uint64_t *solid, int x, int y, int z) { return (int)((solid[(size_t)x * H + y] >> z) & 1); }
static inline uint8_t *lk(orc_links *w, int x, int y, int z) { return &w->link[((size_t)x * H + y) * D + z]; }

static void q_push(orc_links *w, int x, int y, int z) {
	if (w->qtail - w->qhead == w->qcap) { /* grow, keeping order */
		size_t ncap = w->qcap * 2;
		cell_t *nq = (cell_t *)malloc(ncap * sizeof(cell_t));
		for (size_t i = 0; i < w->qcap; i++)
			nq[i] = w->queue[(w->qhead + i) % w->qcap];
		free(w->queue);
		w->queue = nq;
		w->qtail = w->qcap;
		w->qhead = 0;
		w->qcap = ncap;
	}
	cell_t c = {(short)x, (short)y, (short)z};
	w->queue[w->qtail++ % w->qcap] = c;
}
static int q_pop(orc_links *w, cell_t *c) {
	if (w->qhead == w->qtail)
		return 0;
	*c = w->queue[w->qhead++ % w->qcap];
	return 1;
}

orc_links *orc_links_create(void) {
	orc_links *w = (orc_links *)calloc(1, sizeof(*w));
	w->link = (uint8_t *)calloc((size_t)W * H * D, 1);
	w->qcap = 1 << 16;
	w->queue = (cell_t *)malloc(w->qcap * sizeof(cell_t));
	return w;
}
void orc_links_free(orc_links *w) {
	if (!w)
		return;
	free(w->link);
	free(w->queue);
	free(w);
}

/* GameMapWrapper.cpp:50-102 */
void orc_links_rebuild(orc_links *w, const uint64_t *solid) {
	memset(w->link, 0, (size_t)W * H * D);
	w->qhead = w->qtail = 0;
	for (int x = 0; x < W; x++)
		for (int y = 0; y < H; y++)
			*lk(w, x, y, D - 1) = L_ROOT;
	/* the reference's deque starts with 2*W*H default cells (0,0,0) (:62) -- popping them only re-tests the
	 * neighbours of (0,0,0) against links that are still Invalid or not solid, which links (1,0,0) / (0,1,0) /
	 * (0,0,1) to it if they are solid: reproduced by pushing one such cell */
	q_push(w, 0, 0, 0);
	for (int x = 0; x < W; x++)
		for (int y = 0; y < H; y++)
			if (solid_at(solid, x, y, D - 2)) {
				*lk(w, x, y, D - 2) = L_POSZ;
				q_push(w, x, y, D - 2);
			}
	cell_t p;
	while (q_pop(w, &p)) {
		int x = p.x, y = p.y, z = p.z;
		if (x > 0 && solid_at(solid, x - 1, y, z) && *lk(w, x - 1, y, z) == L_INVALID) {
			*lk(w, x - 1, y, z) = L_POSX;
			q_push(w, x - 1, y, z);
		}
		if (x < W - 1 && solid_at(solid, x + 1, y, z) && *lk(w, x + 1, y, z) == L_INVALID) {
			*lk(w, x + 1, y, z) = L_NEGX;
			q_push(w, x + 1, y, z);
		}
		if (y > 0 && solid_at(solid, x, y - 1, z) && *lk(w, x, y - 1, z) == L_INVALID) {
			*lk(w, x, y - 1, z) = L_POSY;
			q_push(w, x, y - 1, z);
		}
		if (y < H - 1 && solid_at(solid, x, y + 1, z) && *lk(w, x, y + 1, z) == L_INVALID) {
			*lk(w, x, y + 1, z) = L_NEGY;
			q_push(w, x, y + 1, z);
		}
		if (z > 0 && solid_at(solid, x, y, z - 1) && *lk(w, x, y, z - 1) == L_INVALID) {
			*lk(w, x, y, z - 1) = L_POSZ;
			q_push(w, x, y, z - 1);
		}
		if (z < D - 1 && solid_at(solid, x, y, z + 1) && *lk(w, x, y, z + 1) == L_INVALID) {
			*lk(w, x, y, z + 1) = L_NEGZ;
			q_push(w, x, y, z + 1);
		}
	}
}

/* the expansion loop shared by AddBlock (:160-192) and RemoveBlocks (:309-345) */
static void relink_from_queue(orc_links *w, const uint64_t *solid) {
	cell_t p;
	while (q_pop(w, &p)) {
		int x = p.x, y = p.y, z = p.z;
		int thisLink = *lk(w, x, y, z);
		if (x > 0 && solid_at(solid, x - 1, y, z) && *lk(w, x - 1, y, z) == L_INVALID && thisLink != L_NEGX) {
			*lk(w, x - 1, y, z) = L_POSX;
			q_push(w, x - 1, y, z);
		}
		if (x < W - 1 && solid_at(solid, x + 1, y, z) && *lk(w, x + 1, y, z) == L_INVALID && thisLink != L_POSX) {
			*lk(w, x + 1, y, z) = L_NEGX;
			q_push(w, x + 1, y, z);
		}
		if (y > 0 && solid_at(solid, x, y - 1, z) && *lk(w, x, y - 1, z) == L_INVALID && thisLink != L_NEGY) {
			*lk(w, x, y - 1, z) = L_POSY;
			q_push(w, x, y - 1, z);
		}
		if (y < H - 1 && solid_at(solid, x, y + 1, z) && *lk(w, x, y + 1, z) == L_INVALID && thisLink != L_POSY) {
			*lk(w, x, y + 1, z) = L_NEGY;
			q_push(w, x, y + 1, z);
		}
		if (z > 0 && solid_at(solid, x, y, z - 1) && *lk(w, x, y, z - 1) == L_INVALID && thisLink != L_NEGZ) {
			*lk(w, x, y, z - 1) = L_POSZ;
			q_push(w, x, y, z - 1);
		}
		if (z < D - 1 && solid_at(solid, x, y, z + 1) && *lk(w, x, y, z + 1) == L_INVALID && thisLink != L_POSZ) {
			*lk(w, x, y, z + 1) = L_NEGZ;
			q_push(w, x, y, z + 1);
		}
	}
}

/* GameMapWrapper.cpp:104-193; GameMap::Set (GameMap.h:104-135) on the caller's arrays */
void orc_links_add_block(orc_links *w, uint64_t *solid, uint32_t *color, int x, int y, int z, uint32_t col) {
	if (*lk(w, x, y, z) != L_INVALID)
		return;
	solid[(size_t)x * H + y] |= 1ull << z;
	color[((size_t)x * H + y) * D + z] = col;
	int l = L_INVALID;
	if (x > 0 && solid_at(solid, x - 1, y, z) && *lk(w, x - 1, y, z) != L_INVALID)
		l = L_NEGX;
	if (x < W - 1 && solid_at(solid, x + 1, y, z) && *lk(w, x + 1, y, z) != L_INVALID)
		l = L_POSX;
	if (y > 0 && solid_at(solid, x, y - 1, z) && *lk(w, x, y - 1, z) != L_INVALID)
		l = L_NEGY;
	if (y < H - 1 && solid_at(solid, x, y + 1, z) && *lk(w, x, y + 1, z) != L_INVALID)
		l = L_POSY;
	if (z > 0 && solid_at(solid, x, y, z - 1) && *lk(w, x, y, z - 1) != L_INVALID)
		l = L_NEGZ;
	if (z < D - 1 && solid_at(solid, x, y, z + 1) && *lk(w, x, y, z + 1) != L_INVALID)
		l = L_POSZ;
	*lk(w, x, y, z) = (uint8_t)l;
	if (l == L_INVALID)
		return;
	w->qhead = w->qtail = 0;
	q_push(w, x, y, z);
	relink_from_queue(w, solid);
}

static inline int eq2(int a, int b, int c, int cond) { return a == b || (cond && a == c); } /* :195-197 */

/* GameMapWrapper.cpp:199-361.  cells: n x {x,y,z}; the cells are cleared in solid (GameMap::Set(.., false, 0)).
 * Returns the number of floating blocks and writes min(count, cap) of them, in the reference's order. */
int orc_links_remove_blocks(orc_links *w, uint64_t *solid, const int16_t *cells, int n, int16_t *out, int cap) {
	if (n <= 0)
		return 0;
	size_t ucap = 1024, un = 0;
	cell_t *unlinked = (cell_t *)malloc(ucap * sizeof(cell_t));
	w->qhead = w->qtail = 0;
	for (int i = 0; i < n; i++) {
		cell_t pos = {cells[i * 3], cells[i * 3 + 1], cells[i * 3 + 2]};
		solid[(size_t)pos.x * H + pos.y] &= ~(1ull << pos.z);
		if (*lk(w, pos.x, pos.y, pos.z) == L_MARKED)
			continue;
		*lk(w, pos.x, pos.y, pos.z) = L_INVALID;
		q_push(w, pos.x, pos.y, pos.z);
		while (q_pop(w, &pos)) {
			if (solid_at(solid, pos.x, pos.y, pos.z)) {
				if (un == ucap) {
					ucap *= 2;
					unlinked = (cell_t *)realloc(unlinked, ucap * sizeof(cell_t));
				}
				unlinked[un++] = pos;
			}
			int x = pos.x, y = pos.y, z = pos.z;
			if (x > 0 && eq2(*lk(w, x - 1, y, z), L_POSX, L_INVALID, solid_at(solid, x - 1, y, z))) {
				*lk(w, x - 1, y, z) = L_MARKED;
				q_push(w, x - 1, y, z);
			}
			if (x < W - 1 && eq2(*lk(w, x + 1, y, z), L_NEGX, L_INVALID, solid_at(solid, x + 1, y, z))) {
				*lk(w, x + 1, y, z) = L_MARKED;
				q_push(w, x + 1, y, z);
			}
			if (y > 0 && eq2(*lk(w, x, y - 1, z), L_POSY, L_INVALID, solid_at(solid, x, y - 1, z))) {
				*lk(w, x, y - 1, z) = L_MARKED;
				q_push(w, x, y - 1, z);
			}
			if (y < H - 1 && eq2(*lk(w, x, y + 1, z), L_NEGY, L_INVALID, solid_at(solid, x, y + 1, z))) {
				*lk(w, x, y + 1, z) = L_MARKED;
				q_push(w, x, y + 1, z);
			}
			if (z > 0 && eq2(*lk(w, x, y, z - 1), L_POSZ, L_INVALID, solid_at(solid, x, y, z - 1))) {
				*lk(w, x, y, z - 1) = L_MARKED;
				q_push(w, x, y, z - 1);
			}
			if (z < D - 1 && eq2(*lk(w, x, y, z + 1), L_NEGZ, L_INVALID, solid_at(solid, x, y, z + 1))) {
				*lk(w, x, y, z + 1) = L_MARKED;
				q_push(w, x, y, z + 1);
			}
		}
	}
	for (size_t i = 0; i < un; i++) {
		cell_t p = unlinked[i];
		if (*lk(w, p.x, p.y, p.z) == L_MARKED)
			*lk(w, p.x, p.y, p.z) = L_INVALID;
	}
	for (size_t i = 0; i < un; i++) {
		cell_t p = unlinked[i];
		int x = p.x, y = p.y, z = p.z;
		if (!solid_at(solid, x, y, z))
			continue;
		int nl = L_INVALID;
		if (z < D - 1 && *lk(w, x, y, z + 1) != L_INVALID)
			nl = L_POSZ;
		else if (x > 0 && *lk(w, x - 1, y, z) != L_INVALID)
			nl = L_NEGX;
		else if (x < W - 1 && *lk(w, x + 1, y, z) != L_INVALID)
			nl = L_POSX;
		else if (y > 0 && *lk(w, x, y - 1, z) != L_INVALID)
			nl = L_NEGY;
		else if (y < H - 1 && *lk(w, x, y + 1, z) != L_INVALID)
			nl = L_POSY;
		else if (z > 0 && *lk(w, x, y, z - 1) != L_INVALID)
			nl = L_NEGZ;
		if (nl != L_INVALID) {
			*lk(w, x, y, z) = (uint8_t)nl;
			q_push(w, x, y, z);
		}
	}
	relink_from_queue(w, solid);
	int nf = 0;
	for (size_t i = 0; i < un; i++) {
		cell_t p = unlinked[i];
		if (!solid_at(solid, p.x, p.y, p.z))
			continue;
		if (*lk(w, p.x, p.y, p.z) == L_INVALID) {
			if (nf < cap) {
				out[nf * 3] = p.x;
				out[nf * 3 + 1] = p.y;
				out[nf * 3 + 2] = p.z;
			}
			nf++;
		}
	}
	free(unlinked);
	return nf;
}

/* link != Invalid per voxel as bit z of out[x][y] */
void orc_links_linked(const orc_links *w, uint64_t *out) {
	for (int x = 0; x < W; x++)
		for (int y = 0; y < H; y++) {
			uint64_t b = 0;
			for (int z = 0; z < D; z++)
				if (w->link[((size_t)x * H + y) * D + z] != L_INVALID)
					b |= 1ull << z;
			out[(size_t)x * H + y] = b;
		}
}

/* Draw/SWFlatMapRenderer.cpp:94-105 GeneratePixel for the whole map: out[y * 512 + x] */
void orc_minimap(const uint64_t *solid, const uint32_t *color, uint32_t *out) {
	for (int y = 0; y < H; y++)
		for (int x = 0; x < W; x++) {
			uint32_t px = 0;
			for (int z = 0; z < D; z++)
				if (solid_at(solid, x, y, z)) {
					uint32_t col = color[((size_t)x * H + y) * D + z];
					col = (col & 0xff00) | ((col & 0xff) << 16) | ((col & 0xff0000) >> 16);
					px = col | 0xff000000u;
					break;
				}
			out[(size_t)y * W + x] = px;
		}
}

/* ---- World::ClusterizeBlocks (Client/World.cpp:154-226) ----
 * The reference puts the cells into an unordered_map<CellPos, bool> (:156-159), then for every still-true entry walks
 * breadth-first (:176-217) over the six neighbours in the order -x +x -y +y -z +z, clearing the flag of each cell it
 * takes (:187).  The outer loop follows the map's iteration order; here it follows the input order, so the pieces are
 * the same and only their numbering differs.  cluster[i] = piece of cell i, returns the number of pieces. */
static int cmp_u32(const void *a, const void *b) {
	const uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b;
	return x < y ? -1 : x > y;
}
static int key_slot(const uint32_t *keys, int n, int x, int y, int z) {
	if (x < 0 || y < 0 || z < 0 || x > 0x7fff || y > 0x7fff || z > 0x7fff)
		return -1;
	const uint32_t k = ((uint32_t)x << 20) | ((uint32_t)y << 8) | (uint32_t)z; /* x < 4096, y < 4096, z < 256 */
	const uint32_t *p = (const uint32_t *)bsearch(&k, keys, (size_t)n, sizeof(uint32_t), cmp_u32);
	return p ? (int)(p - keys) : -1;
}
int orc_clusterize_blocks(const int16_t *cells, int n, int32_t *cluster) {
	if (n <= 0)
		return 0;
	uint32_t *keys = (uint32_t *)malloc((size_t)n * sizeof(uint32_t));
	int m = 0;
	for (int i = 0; i < n; i++)
		keys[i] = ((uint32_t)cells[i * 3] << 20) | ((uint32_t)cells[i * 3 + 1] << 8) | (uint32_t)cells[i * 3 + 2];
	qsort(keys, (size_t)n, sizeof(uint32_t), cmp_u32);
	for (int i = 0; i < n; i++) /* the map holds a cell once (:157-159) */
		if (i == 0 || keys[i] != keys[i - 1])
			keys[m++] = keys[i];
	int32_t *piece = (int32_t *)malloc((size_t)m * sizeof(int32_t)); /* -1 = flag still true */
	int *queue = (int *)malloc((size_t)m * 6 * sizeof(int) + sizeof(int));
	for (int i = 0; i < m; i++)
		piece[i] = -1;
	int count = 0;
	for (int s = 0; s < n; s++) {
		const int start = key_slot(keys, m, cells[s * 3], cells[s * 3 + 1], cells[s * 3 + 2]);
		if (piece[start] >= 0) /* :173-174 */
			continue;
		int head = 0, tail = 0;
		queue[tail++] = start;
		while (head < tail) {
			const int c = queue[head++];
			if (piece[c] >= 0) /* :181-182 */
				continue;
			piece[c] = count; /* :185-187 */
			const int x = (int)(keys[c] >> 20), y = (int)((keys[c] >> 8) & 0xfff), z = (int)(keys[c] & 0xff);
			const int nb[6][3] = {{x - 1, y, z}, {x + 1, y, z}, {x, y - 1, z}, {x, y + 1, z}, {x, y, z - 1}, {x, y, z + 1}};
			for (int k = 0; k < 6; k++) { /* :191-214 */
				const int j = key_slot(keys, m, nb[k][0], nb[k][1], nb[k][2]);
				if (j >= 0 && piece[j] < 0)
					queue[tail++] = j;
			}
		}
		count++; /* :219-221 */
	}
	for (int i = 0; i < n; i++)
		cluster[i] = piece[key_slot(keys, m, cells[i * 3], cells[i * 3 + 1], cells[i * 3 + 2])];
	free(queue);
	free(piece);
	free(keys);
	return count;
}
