// Floating-block detection and the minimap on the resident map.
//   GameMapWrapper::Rebuild       Sources/Client/GameMapWrapper.cpp:50-102   (osb200_links_rebuild)
//   GameMapWrapper::AddBlock      Sources/Client/GameMapWrapper.cpp:104-193  (osb200_links_blocks_added)
//   GameMapWrapper::RemoveBlocks  Sources/Client/GameMapWrapper.cpp:199-361  (osb200_links_remove_blocks)
//   SWFlatMapRenderer::GeneratePixel  Sources/Draw/SWFlatMapRenderer.cpp:94-105 (osb200_minimap)
//
// The reference keeps a spanning forest (one link byte per voxel, 16 MiB) rooted in the bottom layer and repairs it
// with breadth-first walks, one voxel at a time.  What its callers consume is only a set: which solid voxels are no
// longer connected (6-neighbourhood, no wrap) to the ground after a batch of removals.  Here the state is one bit per
// voxel -- linked[x][y] bit z = "connected to the ground", 2 MiB in the map's tiled order -- and connectivity is a
// bit-parallel flood over 64-bit column masks:
//   * inside a column a whole run of solid voxels is filled from any reached voxel with one add (the carry runs
//     through the run) and one bit-reversed add for the other direction: no loop over z;
//   * sideways a column ORs its four neighbours' masks; a CTA iterates a 32 x 32 tile of columns in shared memory to
//     its fixed point, so a launch only has to carry reachability across tile borders; launches repeat until no tile
//     changed (2-4 launches for terrain, more only for long winding tunnels).
// RemoveBlocks = flood from the ground on the edited map, then  fell = linked_before & solid & ~reached  (+ the
// never-linked cluster around a removed voxel that was itself floating, found by a second flood inside the unreached
// voxels).  The result is returned sorted by (x, y, z); the reference returns the same set in breadth-first order.
#include "ctx.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cstring>
#include <vector>

namespace osb {
	constexpr int kLinkTile = 32;
	constexpr uint64_t kBelowRoot = ~(1ull << 63); // z = 63 is Root whatever the map holds and is never expanded

	struct LinksState {
		uint64_t *d_linked = nullptr, *d_reach = nullptr, *d_within = nullptr, *d_reach2 = nullptr; // [tiled] 2 MiB each
		int *d_flags = nullptr; // [0] changed, [1] count, [2] any floating seed
		int *h_flags = nullptr; // pinned
		int16_t *d_cells = nullptr, *d_out = nullptr;
		size_t cellCap = 0, outCap = 0;
		uint32_t *d_minimap = nullptr;
		bool valid = false;
		int lastRounds = 0;
		std::vector<int16_t> lastFloating; // sorted result of the last remove_blocks call (osb200_links_last_floating)
	};

	// all runs of `within` that contain a bit of r (r is a subset of within)
	__device__ __forceinline__ uint64_t vfill(uint64_t within, uint64_t r) {
		const uint64_t up = ((within + r) ^ within) & within;
		const uint64_t wr = __brevll(within), rr = __brevll(r);
		const uint64_t down = __brevll(((wr + rr) ^ wr) & wr);
		return up | down | r;
	}

	// within = solid below the root layer; reach = the seeds of a flood from the ground: every solid voxel of z = 62
	// (they link to Root, :65-70) and the three neighbours of voxel (0,0,0) that Rebuild links through the
	// value-initialised cells its queue starts with (:62)
	// linkedOld == nullptr: Rebuild.  Otherwise (RemoveBlocks) the three quirk cells count as ground only while they
	// still carry the link Rebuild gave them: a voxel ADDED there later went through AddBlock, which knows no such link.
	__global__ void k_links_ground_seeds(const uint64_t *__restrict__ solidT, uint64_t *__restrict__ within,
	                                     uint64_t *__restrict__ reach, const uint64_t *__restrict__ linkedOld) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= OSB200_MAP_W * OSB200_MAP_H)
			return;
		const uint64_t w = solidT[i] & kBelowRoot;
		const uint64_t quirk = linkedOld ? w & linkedOld[i] : w;
		uint64_t r = w & (1ull << 62);
		if (i == (int)tiled_index(1, 0) || i == (int)tiled_index(0, 1))
			r |= quirk & 1ull;
		if (i == (int)tiled_index(0, 0))
			r |= quirk & 2ull;
		within[i] = w;
		reach[i] = vfill(w, r);
	}

	// AddBlock (:104-193): seeds of the incremental flood = everything that is linked already (and still solid) plus
	// the solid voxels of z = 62, which link to Root below them (:141-144).  Rebuild's (0,0,0) quirk is NOT re-applied:
	// AddBlock has no such path, a voxel added at (1,0,0) / (0,1,0) / (0,0,1) next to nothing linked stays Invalid.
	__global__ void k_links_add_seeds(const uint64_t *__restrict__ solidT, const uint64_t *__restrict__ linked,
	                                  uint64_t *__restrict__ within, uint64_t *__restrict__ reach) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= OSB200_MAP_W * OSB200_MAP_H)
			return;
		const uint64_t w = solidT[i] & kBelowRoot;
		within[i] = w;
		reach[i] = vfill(w, (linked[i] & w) | (w & (1ull << 62)));
	}

	__global__ void __launch_bounds__(kLinkTile *kLinkTile)
	  k_links_flood(const uint64_t *__restrict__ within, uint64_t *__restrict__ reach, int *__restrict__ changed) {
		__shared__ uint64_t sR[kLinkTile + 2][kLinkTile + 2]; // [x][y] with a one-column halo
		const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
		// a warp covers two 4x4 storage tiles that are adjacent in y: 256 contiguous bytes
		const int lx = (warp >> 2) * 4 + ((lane >> 2) & 3), ly = (warp & 3) * 8 + ((lane & 3) | ((lane >> 4) << 2));
		const int x0 = blockIdx.x * kLinkTile, y0 = blockIdx.y * kLinkTile;
		const uint32_t t = tiled_index(x0 + lx, y0 + ly);
		const uint64_t W = within[t];
		uint64_t r = reach[t];
		const uint64_t r0 = r;
		sR[lx + 1][ly + 1] = r;
		if (threadIdx.x < 4 * kLinkTile) { // halo: the map edge has no neighbour (no wrap, :76-91)
			const int side = threadIdx.x >> 5, k = threadIdx.x & 31;
			const int hx = side == 0 ? x0 - 1 : (side == 1 ? x0 + kLinkTile : x0 + k);
			const int hy = side == 2 ? y0 - 1 : (side == 3 ? y0 + kLinkTile : y0 + k);
			uint64_t v = 0;
			if (hx >= 0 && hx < OSB200_MAP_W && hy >= 0 && hy < OSB200_MAP_H)
				v = reach[tiled_index(hx, hy)];
			sR[hx - x0 + 1][hy - y0 + 1] = v;
		}
		__syncthreads();
		for (int it = 0; it < 4 * kLinkTile; it++) {
			const uint64_t nb = sR[lx][ly + 1] | sR[lx + 2][ly + 1] | sR[lx + 1][ly] | sR[lx + 1][ly + 2];
			const uint64_t nr = vfill(W, r | (nb & W));
			const bool ch = nr != r;
			r = nr;
			__syncthreads();
			if (ch)
				sR[lx + 1][ly + 1] = r;
			if (!__syncthreads_or(ch))
				break;
		}
		if (r != r0) {
			reach[t] = r;
			*changed = 1;
		}
	}

	// a removed voxel that was not linked drags its whole never-linked cluster with it (the unlink walk follows
	// Invalid && solid neighbours, :236-265): seed the second flood with its unreached solid neighbours
	__global__ void k_links_float_seeds(const int16_t *__restrict__ cells, int n, const uint64_t *__restrict__ linkedOld,
	                                    const uint64_t *__restrict__ within, const uint64_t *__restrict__ reach,
	                                    uint64_t *__restrict__ seeds, int *__restrict__ flags) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n)
			return;
		const int x = cells[i * 3], y = cells[i * 3 + 1], z = cells[i * 3 + 2];
		if ((linkedOld[tiled_index(x, y)] >> z) & 1ull)
			return;
		auto seed = [&](int nx, int ny, int nz) {
			if (nx < 0 || nx >= OSB200_MAP_W || ny < 0 || ny >= OSB200_MAP_H || nz < 0 || nz > 62)
				return;
			const uint32_t t = tiled_index(nx, ny);
			const uint64_t bit = 1ull << nz;
			if (within[t] & ~reach[t] & bit) {
				atomicOr(reinterpret_cast<unsigned long long *>(seeds + t), (unsigned long long)bit);
				flags[2] = 1;
			}
		};
		seed(x - 1, y, z);
		seed(x + 1, y, z);
		seed(x, y - 1, z);
		seed(x, y + 1, z);
		seed(x, y, z - 1);
		seed(x, y, z + 1);
	}
	// second flood runs inside the voxels the ground flood did not reach
	__global__ void k_links_float_prepare(const uint64_t *__restrict__ within, const uint64_t *__restrict__ reach,
	                                      uint64_t *__restrict__ within2, uint64_t *__restrict__ seeds) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= OSB200_MAP_W * OSB200_MAP_H)
			return;
		const uint64_t w2 = within[i] & ~reach[i];
		within2[i] = w2;
		seeds[i] = vfill(w2, seeds[i] & w2);
	}

	// fell = (linked before & solid & not reached) | floating cluster; one thread per column, any order (sorted on the host)
	__global__ void k_links_emit(const uint64_t *__restrict__ linkedOld, const uint64_t *__restrict__ within,
	                             const uint64_t *__restrict__ reach, const uint64_t *__restrict__ cluster,
	                             int16_t *__restrict__ out, int cap, int *__restrict__ flags) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= OSB200_MAP_W * OSB200_MAP_H)
			return;
		const int x = i >> 9, y = i & 511;
		const uint32_t t = tiled_index(x, y);
		uint64_t m = linkedOld[t] & within[t] & ~reach[t];
		if (cluster)
			m |= cluster[t];
		while (m) {
			const int z = __ffsll((long long)m) - 1;
			m &= m - 1;
			const int k = atomicAdd(&flags[1], 1);
			if (out && k < cap) {
				out[k * 3] = (int16_t)x;
				out[k * 3 + 1] = (int16_t)y;
				out[k * 3 + 2] = (int16_t)z;
			}
		}
	}

	__global__ void k_links_finish(const uint64_t *__restrict__ reach, uint64_t *__restrict__ linked) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i < OSB200_MAP_W * OSB200_MAP_H)
			linked[i] = reach[i] | (1ull << 63); // Root layer (:56-58)
	}
	__global__ void k_links_untile(const uint64_t *__restrict__ linkedT, uint64_t *__restrict__ outXY) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i < OSB200_MAP_W * OSB200_MAP_H)
			outXY[i] = linkedT[tiled_index(i >> 9, i & 511)];
	}

	// SWFlatMapRenderer::GeneratePixel (:94-105): the colour of the topmost solid voxel, R and B swapped, opaque
	__global__ void k_minimap(const uint64_t *__restrict__ solidT, const uint32_t *__restrict__ colorT,
	                          uint32_t *__restrict__ out) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x; // i = y * 512 + x
		if (i >= OSB200_MAP_W * OSB200_MAP_H)
			return;
		const int x = i & 511, y = i >> 9;
		const uint32_t t = tiled_index(x, y);
		const uint64_t s = solidT[t];
		uint32_t px = 0;
		if (s) {
			const int z = __ffsll((long long)s) - 1;
			const uint32_t col = colorT[(size_t)t * 64 + z];
			px = (col & 0xff00u) | ((col & 0xffu) << 16) | ((col & 0xff0000u) >> 16) | 0xff000000u;
		}
		out[i] = px;
	}

	// ---- World::ClusterizeBlocks (Sources/Client/World.cpp:154-226): 6-connected components of a set of cells ----
	// The reference walks the set breadth-first through an unordered_map, one cell at a time.  Here the cells arrive
	// become 24-bit keys (x << 15 | y << 6 | z, the order of CellPos::operator<) sorted on the device (cub radix sort,
	// three 8-bit passes), a thread per cell looks its +x, +y and +z neighbours up by binary search and joins them in a lock-free union-find whose roots are the smallest
	// index of a component (a larger root is only ever hooked below a smaller one, with atomicCAS), so the labels do
	// not depend on the interleaving of the threads.
	__device__ __forceinline__ int cl_find(int *parent, int i) {
		for (;;) {
			const int p = ((volatile int *)parent)[i];
			if (p == i)
				return i;
			const int g = ((volatile int *)parent)[p];
			if (g != p)
				atomicMin(&parent[i], g); // path halving; parents only move towards smaller indices
			i = p;
		}
	}
	__device__ __forceinline__ void cl_union(int *parent, int a, int b) {
		for (;;) {
			a = cl_find(parent, a);
			b = cl_find(parent, b);
			if (a == b)
				return;
			const int hi = max(a, b), lo = min(a, b);
			if (atomicCAS(&parent[hi], hi, lo) == hi)
				return;
		}
	}
	__device__ __forceinline__ int cl_lookup(const uint32_t *__restrict__ keys, int n, uint32_t k) {
		int lo = 0, hi = n;
		while (lo < hi) {
			const int mid = (lo + hi) >> 1;
			if (__ldg(&keys[mid]) < k)
				lo = mid + 1;
			else
				hi = mid;
		}
		return lo < n && __ldg(&keys[lo]) == k ? lo : -1;
	}
	__global__ void k_cluster_init(int *__restrict__ parent, int n) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i < n)
			parent[i] = i;
	}
	__global__ void k_cluster_union(const uint32_t *__restrict__ keys, int n, int *parent) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n)
			return;
		const uint32_t k = keys[i];
		if (i > 0 && keys[i - 1] == k) // the same cell twice is one cell (the reference's map holds it once)
			cl_union(parent, i, i - 1);
		int j;
		if ((k & 63u) != 63u && i + 1 < n && keys[i + 1] == k + 1) // +z is the next key when it is there
			cl_union(parent, i, i + 1);
		if (((k >> 6) & 511u) != 511u && (j = cl_lookup(keys, n, k + (1u << 6))) >= 0)
			cl_union(parent, i, j);
		if ((k >> 15) != 511u && (j = cl_lookup(keys, n, k + (1u << 15))) >= 0)
			cl_union(parent, i, j);
	}
	__global__ void k_cluster_label(int *parent, int n, int *__restrict__ isRoot) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i < n) {
			const int r = cl_find(parent, i); // roots are fixed by now: only pointers below them are rewritten
			parent[i] = r;
			isRoot[i] = r == i;
		}
	}
	// cells -> 24-bit keys + the slot they came from; flags[0] = smallest index of a cell outside the map
	__global__ void k_cluster_keys(const int16_t *__restrict__ cells, int n, uint32_t *__restrict__ keys,
	                               uint32_t *__restrict__ slot, int *__restrict__ flags) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n)
			return;
		const int x = cells[i * 3], y = cells[i * 3 + 1], z = cells[i * 3 + 2];
		const bool ok = x >= 0 && x < OSB200_MAP_W && y >= 0 && y < OSB200_MAP_H && z >= 0 && z < OSB200_MAP_D;
		if (!ok)
			atomicMin(&flags[0], i);
		keys[i] = ok ? ((uint32_t)x << 15 | (uint32_t)y << 6 | (uint32_t)z) : 0u;
		slot[i] = (uint32_t)i;
	}
	// label[input cell] = number of its root; flags[1] = number of pieces
	__global__ void k_cluster_scatter(const int *__restrict__ root, const int *__restrict__ number,
	                                  const int *__restrict__ isRoot, const uint32_t *__restrict__ slot, int n,
	                                  int *__restrict__ label, int *__restrict__ count) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n)
			return;
		label[slot[i]] = number[root[i]];
		if (i == n - 1)
			*count = number[i] + isRoot[i];
	}

	static int links_init(osb200_ctx *c) {
		if (c->links)
			return 0;
		auto *S = new LinksState();
		c->links = S;
		const size_t nb = (size_t)OSB200_MAP_W * OSB200_MAP_H * 8;
		OSB_CUDA(cudaMalloc(&S->d_linked, nb));
		OSB_CUDA(cudaMalloc(&S->d_reach, nb));
		OSB_CUDA(cudaMalloc(&S->d_within, nb));
		OSB_CUDA(cudaMalloc(&S->d_reach2, nb));
		OSB_CUDA(cudaMalloc(&S->d_flags, 4 * sizeof(int)));
		OSB_CUDA(cudaMallocHost(&S->h_flags, 4 * sizeof(int)));
		OSB_CUDA(cudaMalloc(&S->d_minimap, (size_t)OSB200_MAP_W * OSB200_MAP_H * 4));
		return 0;
	}
	void links_destroy(osb200_ctx *c) {
		LinksState *S = c->links;
		if (!S)
			return;
		cudaFree(S->d_linked);
		cudaFree(S->d_reach);
		cudaFree(S->d_within);
		cudaFree(S->d_reach2);
		cudaFree(S->d_flags);
		cudaFreeHost(S->h_flags);
		cudaFree(S->d_cells);
		cudaFree(S->d_out);
		cudaFree(S->d_minimap);
		delete S;
		c->links = nullptr;
	}
	void links_map_replaced(osb200_ctx *c) {
		if (c->links)
			c->links->valid = false;
	}

	// repeat the tile flood until no tile changed
	static int flood(osb200_ctx *c, const uint64_t *within, uint64_t *reach) {
		LinksState *S = c->links;
		const dim3 grid(OSB200_MAP_W / kLinkTile, OSB200_MAP_H / kLinkTile);
		int rounds = 0;
		for (;;) {
			OSB_CUDA(cudaMemsetAsync(S->d_flags, 0, sizeof(int), c->stream));
			{
				KLaunch kl(c, K_LINKS_FLOOD);
				k_links_flood<<<grid, kLinkTile * kLinkTile, 0, c->stream>>>(within, reach, S->d_flags);
			}
			OSB_CUDA(cudaMemcpyAsync(S->h_flags, S->d_flags, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
			OSB_CUDA(cudaStreamSynchronize(c->stream));
			rounds++;
			if (!S->h_flags[0])
				break;
			if (rounds > 4096) {
				set_error("links: flood did not converge");
				return -2;
			}
		}
		S->lastRounds += rounds;
		return 0;
	}
	static int ground_flood(osb200_ctx *c, bool rebuild) {
		LinksState *S = c->links;
		const int n = OSB200_MAP_W * OSB200_MAP_H;
		k_links_ground_seeds<<<n / 256, 256, 0, c->stream>>>(c->d_solid, S->d_within, S->d_reach, rebuild ? nullptr : S->d_linked);
		c->launches++;
		return flood(c, S->d_within, S->d_reach);
	}
	static int links_ready(osb200_ctx *c, const char *who, bool needValid) {
		if (!c) {
			set_error("%s: null context", who);
			return -1;
		}
		if (!c->mapLoaded) {
			set_error("%s: no map uploaded", who);
			return -1;
		}
		if (c->anyDirty) {
			set_error("%s: unflushed edits (call osb200_map_flush first)", who);
			return -1;
		}
		if (needValid && (!c->links || !c->links->valid)) {
			set_error("%s: link state not built for this map (call osb200_links_rebuild)", who);
			return -1;
		}
		return 0;
	}
} // namespace osb

extern "C" {

int osb200_links_rebuild(osb200_ctx *c) {
	using namespace osb;
	if (int rc = links_ready(c, "osb200_links_rebuild", false))
		return rc;
	OSB_CUDA(cudaSetDevice(c->device));
	if (links_init(c))
		return -2;
	LinksState *S = c->links;
	S->lastRounds = 0;
	if (int rc = ground_flood(c, true))
		return rc;
	k_links_finish<<<OSB200_MAP_W * OSB200_MAP_H / 256, 256, 0, c->stream>>>(S->d_reach, S->d_linked);
	c->launches++;
	OSB_CUDA(cudaGetLastError());
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	S->valid = true;
	return S->lastRounds;
}

int osb200_links_blocks_added(osb200_ctx *c) {
	// AddBlock links a new voxel next to a linked one and walks on into never-linked neighbours (:127-192): after any
	// number of additions the linked set is again "connected to the ground"
	// The linked set is closed under solid adjacency (Rebuild and RemoveBlocks leave it so), hence flooding from the
	// whole linked set through the solid voxels reaches exactly what the per-voxel walks of AddBlock would link: new
	// voxels that touch a linked one, and on through never-linked neighbours.  Usually one flood launch (nothing moves
	// outside the tiles that hold the new voxels).
	using namespace osb;
	if (int rc = links_ready(c, "osb200_links_blocks_added", true))
		return rc;
	OSB_CUDA(cudaSetDevice(c->device));
	LinksState *S = c->links;
	S->lastRounds = 0;
	const int nCols = OSB200_MAP_W * OSB200_MAP_H;
	k_links_add_seeds<<<nCols / 256, 256, 0, c->stream>>>(c->d_solid, S->d_linked, S->d_within, S->d_reach);
	c->launches++;
	if (int rc = flood(c, S->d_within, S->d_reach))
		return rc;
	k_links_finish<<<nCols / 256, 256, 0, c->stream>>>(S->d_reach, S->d_linked);
	c->launches++;
	OSB_CUDA(cudaGetLastError());
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return S->lastRounds;
}

int osb200_links_remove_blocks(osb200_ctx *c, const int16_t *cells, int n, int16_t *floating, int cap) {
	using namespace osb;
	if (int rc = links_ready(c, "osb200_links_remove_blocks", true))
		return rc;
	if (n < 0 || (n && !cells) || cap < 0 || (cap && !floating)) {
		set_error("osb200_links_remove_blocks: bad argument");
		return -1;
	}
	if (n == 0)
		return 0; // :202-203
	for (int i = 0; i < n; i++) {
		const int x = cells[i * 3], y = cells[i * 3 + 1], z = cells[i * 3 + 2];
		if (x < 0 || x >= OSB200_MAP_W || y < 0 || y >= OSB200_MAP_H || z < 0 || z >= 63) {
			set_error("osb200_links_remove_blocks: cell %d out of range (the bottom layer cannot be removed)", i);
			return -1;
		}
	}
	OSB_CUDA(cudaSetDevice(c->device));
	LinksState *S = c->links;
	S->lastRounds = 0;
	const int nCols = OSB200_MAP_W * OSB200_MAP_H;
	if ((size_t)n > S->cellCap) {
		cudaFree(S->d_cells);
		S->d_cells = nullptr;
		S->cellCap = 0;
		OSB_CUDA(cudaMalloc(&S->d_cells, ((size_t)n + 1024) * 3 * sizeof(int16_t)));
		S->cellCap = (size_t)n + 1024;
	}
	OSB_CUDA(cudaMemcpyAsync(S->d_cells, cells, (size_t)n * 3 * sizeof(int16_t), cudaMemcpyHostToDevice, c->stream));
	if (int rc = ground_flood(c, false))
		return rc;
	OSB_CUDA(cudaMemsetAsync(S->d_flags, 0, 4 * sizeof(int), c->stream));
	OSB_CUDA(cudaMemsetAsync(S->d_reach2, 0, (size_t)nCols * 8, c->stream));
	k_links_float_seeds<<<(n + 127) / 128, 128, 0, c->stream>>>(S->d_cells, n, S->d_linked, S->d_within, S->d_reach,
	                                                          S->d_reach2, S->d_flags);
	c->launches++;
	OSB_CUDA(cudaMemcpyAsync(S->h_flags, S->d_flags, 4 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	const bool haveCluster = S->h_flags[2] != 0;
	uint64_t *d_within2 = nullptr;
	if (haveCluster) {
		if (c->ensure_scratch((size_t)nCols * 8))
			return -2;
		d_within2 = reinterpret_cast<uint64_t *>(c->d_scratch);
		k_links_float_prepare<<<nCols / 256, 256, 0, c->stream>>>(S->d_within, S->d_reach, d_within2, S->d_reach2);
		c->launches++;
		if (int rc = flood(c, d_within2, S->d_reach2))
			return rc;
	}
	// count first, then emit into a buffer that is large enough
	OSB_CUDA(cudaMemsetAsync(S->d_flags, 0, 4 * sizeof(int), c->stream));
	k_links_emit<<<nCols / 256, 256, 0, c->stream>>>(S->d_linked, S->d_within, S->d_reach, haveCluster ? S->d_reach2 : nullptr,
	                                               nullptr, 0, S->d_flags);
	c->launches++;
	OSB_CUDA(cudaMemcpyAsync(S->h_flags, S->d_flags, 4 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	const int count = S->h_flags[1];
	if ((size_t)count > S->outCap) {
		cudaFree(S->d_out);
		S->d_out = nullptr;
		S->outCap = 0;
		const size_t want = (size_t)count + ((size_t)count >> 1) + 4096;
		OSB_CUDA(cudaMalloc(&S->d_out, want * 3 * sizeof(int16_t)));
		S->outCap = want;
	}
	if (count) {
		OSB_CUDA(cudaMemsetAsync(S->d_flags, 0, 4 * sizeof(int), c->stream));
		k_links_emit<<<nCols / 256, 256, 0, c->stream>>>(S->d_linked, S->d_within, S->d_reach,
		                                               haveCluster ? S->d_reach2 : nullptr, S->d_out, (int)S->outCap, S->d_flags);
		c->launches++;
	}
	k_links_finish<<<nCols / 256, 256, 0, c->stream>>>(S->d_reach, S->d_linked);
	c->launches++;
	OSB_CUDA(cudaGetLastError());
	std::vector<int16_t> host((size_t)count * 3);
	if (count)
		OSB_CUDA(cudaMemcpyAsync(host.data(), S->d_out, (size_t)count * 3 * sizeof(int16_t), cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	// deterministic order: (x, y, z) ascending, CellPos::operator< (GameMapWrapper.h:35-48)
	std::vector<int> idx(count);
	for (int i = 0; i < count; i++)
		idx[i] = i;
	std::sort(idx.begin(), idx.end(), [&](int a, int b) {
		return std::lexicographical_compare(&host[(size_t)a * 3], &host[(size_t)a * 3] + 3, &host[(size_t)b * 3], &host[(size_t)b * 3] + 3);
	});
	S->lastFloating.resize((size_t)count * 3);
	for (int i = 0; i < count; i++)
		std::memcpy(&S->lastFloating[(size_t)i * 3], &host[(size_t)idx[i] * 3], 3 * sizeof(int16_t));
	if (cap > 0 && count > 0)
		std::memcpy(floating, S->lastFloating.data(), (size_t)std::min(count, cap) * 3 * sizeof(int16_t));
	return count;
}

int osb200_links_last_floating(osb200_ctx *c, int16_t *floating, int cap) {
	if (!c || !c->links || cap < 0 || (cap && !floating)) {
		osb::set_error("osb200_links_last_floating: bad argument");
		return -1;
	}
	const int count = (int)(c->links->lastFloating.size() / 3);
	if (cap > 0 && count > 0)
		std::memcpy(floating, c->links->lastFloating.data(), (size_t)std::min(count, cap) * 3 * sizeof(int16_t));
	return count;
}

int osb200_links_read(osb200_ctx *c, uint64_t *out) {
	using namespace osb;
	if (int rc = links_ready(c, "osb200_links_read", true))
		return rc;
	if (!out) {
		set_error("osb200_links_read: null argument");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	const int nCols = OSB200_MAP_W * OSB200_MAP_H;
	if (c->ensure_scratch((size_t)nCols * 8))
		return -2;
	auto *tmp = reinterpret_cast<uint64_t *>(c->d_scratch);
	k_links_untile<<<nCols / 256, 256, 0, c->stream>>>(c->links->d_linked, tmp);
	c->launches++;
	OSB_CUDA(cudaMemcpyAsync(out, tmp, (size_t)nCols * 8, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_clusterize_blocks(osb200_ctx *c, const int16_t *cells, int n, int32_t *cluster) {
	using namespace osb;
	if (!c || n < 0 || (n && (!cells || !cluster))) {
		set_error("osb200_clusterize_blocks: bad argument");
		return -1;
	}
	if (n == 0)
		return 0;
	OSB_CUDA(cudaSetDevice(c->device));
	cudaStream_t st = c->stream;
	// scratch: cells | keys x2 | slots x2 (which input cell a sorted slot came from) | parent | number | labels | flags | cub
	const size_t N = ((size_t)n + 63) & ~(size_t)63;
	size_t sortBytes = 0, scanBytes = 0;
	OSB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sortBytes, (const uint32_t *)nullptr, (uint32_t *)nullptr,
	                                         (const uint32_t *)nullptr, (uint32_t *)nullptr, n, 0, 24, st));
	OSB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scanBytes, (const int *)nullptr, (int *)nullptr, n, st));
	const size_t tempBytes = (std::max(sortBytes, scanBytes) + 255) & ~(size_t)255;
	if (c->ensure_scratch(N * 8 + N * 4 * 8 + 256 + tempBytes))
		return -2;
	uint8_t *base = reinterpret_cast<uint8_t *>(c->d_scratch);
	int16_t *d_cells = reinterpret_cast<int16_t *>(base);
	uint32_t *d_keysIn = reinterpret_cast<uint32_t *>(base + N * 8);
	uint32_t *d_keys = d_keysIn + N, *d_slotIn = d_keys + N, *d_slot = d_slotIn + N;
	int *d_parent = reinterpret_cast<int *>(d_slot + N), *d_number = d_parent + N, *d_label = d_number + N;
	int *d_isRoot = d_label + N;
	int *d_flags = d_isRoot + N; // [0] smallest index of a cell outside the map, [1] number of pieces
	void *d_temp = reinterpret_cast<uint8_t *>(d_flags) + 256;
	OSB_CUDA(cudaMemcpyAsync(d_cells, cells, (size_t)n * 3 * sizeof(int16_t), cudaMemcpyHostToDevice, st));
	OSB_CUDA(cudaMemsetAsync(d_flags, 0x7f, sizeof(int), st));
	const int grid = (n + 255) / 256;
	k_cluster_keys<<<grid, 256, 0, st>>>(d_cells, n, d_keysIn, d_slotIn, d_flags);
	size_t tb = tempBytes;
	OSB_CUDA(cub::DeviceRadixSort::SortPairs(d_temp, tb, d_keysIn, d_keys, d_slotIn, d_slot, n, 0, 24, st));
	k_cluster_init<<<grid, 256, 0, st>>>(d_parent, n);
	{
		KLaunch kl(c, K_CLUSTER_UNION);
		k_cluster_union<<<grid, 256, 0, st>>>(d_keys, n, d_parent);
	}
	k_cluster_label<<<grid, 256, 0, st>>>(d_parent, n, d_isRoot);
	// a root is the smallest sorted slot of its piece: numbering the roots in slot order numbers the pieces by their
	// smallest cell
	tb = tempBytes;
	OSB_CUDA(cub::DeviceScan::ExclusiveSum(d_temp, tb, d_isRoot, d_number, n, st));
	k_cluster_scatter<<<grid, 256, 0, st>>>(d_parent, d_number, d_isRoot, d_slot, n, d_label, d_flags + 1);
	c->launches += 4;
	OSB_CUDA(cudaGetLastError());
	int flags[2] = {0, 0};
	OSB_CUDA(cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, st));
	OSB_CUDA(cudaStreamSynchronize(st));
	if (flags[0] < n) {
		set_error("osb200_clusterize_blocks: cell %d outside the map", flags[0]);
		return -1;
	}
	OSB_CUDA(cudaMemcpyAsync(cluster, d_label, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
	OSB_CUDA(cudaStreamSynchronize(st));
	return flags[1];
}

int osb200_links_last_rounds(osb200_ctx *c) { return c && c->links ? c->links->lastRounds : -1; }

int osb200_minimap(osb200_ctx *c, uint32_t *out, const uint32_t **device) {
	using namespace osb;
	if (int rc = links_ready(c, "osb200_minimap", false))
		return rc;
	OSB_CUDA(cudaSetDevice(c->device));
	if (links_init(c))
		return -2;
	LinksState *S = c->links;
	const int nCols = OSB200_MAP_W * OSB200_MAP_H;
	{
		KLaunch kl(c, K_MINIMAP);
		k_minimap<<<nCols / 256, 256, 0, c->stream>>>(c->d_solid, c->d_color, S->d_minimap);
	}
	OSB_CUDA(cudaGetLastError());
	if (out)
		OSB_CUDA(cudaMemcpyAsync(out, S->d_minimap, (size_t)nCols * 4, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	if (device)
		*device = S->d_minimap;
	return 0;
}

} // extern "C"
