// VXL map format on the device: GameMap::Load (Sources/Client/GameMap.cpp:519-596) straight into the resident tiled
// map, and GameMap::Save (:93-160, IsSurface :65-83) out of it.
//
// The file is a y-major sequence of 262 144 columns, each a chain of variable-length span records, with no index:
// where a column starts is only known after walking every span before it.  Design:
//   1. k_vxl_jump     every 4-byte position of the file is treated as a possible column start; one thread walks the
//                     span chain from it (<= 64 headers) and records where the next column would start.  Positions in
//                     the middle of colour data yield garbage, bounded walks -- they are never followed.
//   2. k_vxl_double   pointer doubling: in round j the first 2^j known column starts produce the next 2^j through
//                     jump^(2^j), and the jump table is squared.  18 rounds give all 2^18 column starts.
//   3. k_vxl_decode   one thread per column replays the reference's span loop (same write order, including the
//                     "z = 63 copies the colour of z = 62" rule) into the tiled solid / colour arrays.
// Save: k_vxl_measure (bytes per column, from the 64-bit column masks: IsSurface is five mask operations),
// a device-wide exclusive scan, k_vxl_emit (one thread per column writes its records).
#include "ctx.h"

#include <cstring>
#include <vector>

namespace osb {
	constexpr int kCols = OSB200_MAP_W * OSB200_MAP_H;
	constexpr uint32_t kBad = 0xffffffffu; // "no column starts here / walk left the file"

	struct VxlState {
		uint8_t *d_file = nullptr;
		size_t fileCap = 0;
		uint32_t *d_jumpA = nullptr, *d_jumpB = nullptr; // nWords + 1 entries each
		size_t jumpCap = 0;
		uint32_t *d_start = nullptr; // kCols + 1 (column starts in words; [kCols] = end of the last column)
		uint32_t *d_sizes = nullptr; // kCols + 1 (Save: bytes per column -> exclusive scan)
		uint32_t *d_blockSums = nullptr;
		int *d_err = nullptr;
	};

	__device__ __forceinline__ int s8(const uint8_t *f, size_t p) { return (int)(int8_t)f[p]; }

	// Where does the column that starts at word i end?  (the `pos` arithmetic of the load loop, :533-585)
	__global__ void k_vxl_jump(const uint8_t *__restrict__ file, uint32_t nWords, uint32_t *__restrict__ jump) {
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i > nWords)
			return;
		if (i == nWords) {
			jump[i] = kBad; // the end of the file starts nothing
			return;
		}
		long long p = i; // in words
		uint32_t out = kBad;
		for (int guard = 0; guard < 64; guard++) {
			if (p < 0 || p >= (long long)nWords)
				break;
			const int chunks = s8(file, (size_t)p * 4);
			if (chunks == 0) {
				const int topStart = s8(file, (size_t)p * 4 + 1), topEnd = s8(file, (size_t)p * 4 + 2);
				const long long e = p + (topEnd - topStart + 1) + 1;
				if (e > p && e <= (long long)nWords)
					out = (uint32_t)e;
				break;
			}
			if (chunks < 0)
				break; // the reference would walk backwards; never on a real map
			p += chunks;
		}
		jump[i] = out;
	}

	// Round j of the doubling: start[k + have] = jump^(have)[start[k]] for k < have; jumpOut = jumpIn o jumpIn.
	__global__ void k_vxl_double(const uint32_t *__restrict__ jumpIn, uint32_t *__restrict__ jumpOut, uint32_t nWords,
	                             uint32_t *__restrict__ start, uint32_t have, uint32_t want) {
		const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i <= nWords) {
			const uint32_t a = jumpIn[i];
			jumpOut[i] = a == kBad ? kBad : jumpIn[a];
		}
		if (i < have && i + have < want) {
			const uint32_t s = start[i];
			start[i + have] = s == kBad ? kBad : jumpIn[s];
		}
	}

	__device__ __forceinline__ uint32_t file_color(const uint8_t *f, size_t p) { // swapColor, GameMap.cpp:509-517
		return (uint32_t)f[p + 2] | ((uint32_t)f[p + 1] << 8) | ((uint32_t)f[p] << 16) | (100u << 24);
	}

	// The column loop of GameMap::Load for column k = x + 512 y, writing the tiled resident arrays.  The colour array
	// was zeroed beforehand (a fresh GameMap's colorMap).
	__global__ void k_vxl_decode(const uint8_t *__restrict__ file, uint32_t nWords, const uint32_t *__restrict__ start,
	                             uint64_t *__restrict__ solidT, uint32_t *__restrict__ colorT, int *__restrict__ err) {
		const int k = blockIdx.x * blockDim.x + threadIdx.x;
		if (k >= kCols)
			return;
		const int x = k & 511, y = k >> 9;
		const uint32_t s0 = start[k], s1 = start[k + 1];
		if (s0 == kBad || s1 == kBad || s0 >= nWords || s1 > nWords || s1 <= s0) {
			atomicExch(err, 1); // truncated or malformed input (never read the stream outside [0, nWords) words)
			return;
		}
		const uint32_t t = tiled_index(x, y);
		uint32_t *col = colorT + (size_t)t * 64;
		uint64_t bits = ~0ull;
		uint32_t c62 = 0; // colour currently stored at z = 62
		bool bad = false;
		auto put = [&](int z, uint32_t c) {
			if (z < 0 || z > 63) { // GameMap::Set asserts the range
				bad = true;
				return;
			}
			col[z] = c;
			bits |= 1ull << z;
			if (z == 62)
				c62 = c;
		};
		size_t pos = (size_t)s0 * 4;
		const size_t end = (size_t)s1 * 4;
		int airFrom = 0;
		for (int guard = 0; guard < 64 && !bad; guard++) {
			const int chunks = s8(file, pos), topStart = s8(file, pos + 1), topEnd = s8(file, pos + 2);
			for (int z = airFrom; z < topStart; z++) {
				if (z < 0 || z > 63) {
					bad = true;
					break;
				}
				bits &= ~(1ull << z);
			}
			size_t cpos = pos + 4;
			for (int z = topStart; z <= topEnd && !bad; z++, cpos += 4) {
				if (cpos + 4 > end) { // colours never run past the column on a well-formed map
					bad = true;
					break;
				}
				put(z, file_color(file, cpos));
			}
			if (topEnd == 62)
				put(63, c62);
			const int topLen = topEnd - topStart + 1;
			if (chunks == 0)
				break;
			const int bottomLen = (chunks - 1) - topLen;
			pos += (size_t)chunks * 4;
			if (pos + 4 > end) { // the walk in k_vxl_jump guarantees this for every header it accepted
				bad = true;
				break;
			}
			const int bottomEnd = s8(file, pos + 3);
			for (int z = bottomEnd - bottomLen; z < bottomEnd && !bad; z++, cpos += 4) {
				if (cpos + 4 > pos) { // colours of a span never run into the next header on a well-formed map
					bad = true;
					break;
				}
				put(z, file_color(file, cpos));
			}
			if (bottomEnd == 63)
				put(63, c62);
			airFrom = bottomEnd;
		}
		if (bad)
			atomicExch(err, 1);
		solidT[t] = bits;
	}

	// tiled -> [x][y] / [x][y][z] (the host arrays of GameMap / VoxelMap); 16 threads per column
	__global__ void k_map_untile(const uint64_t *__restrict__ solidT, const uint32_t *__restrict__ colorT,
	                             uint64_t *__restrict__ solidXY, uint32_t *__restrict__ colorXYZ) {
		const int col = blockIdx.x * (blockDim.x >> 4) + (threadIdx.x >> 4);
		const int part = threadIdx.x & 15;
		if (col >= kCols)
			return;
		const int x = col >> 9, y = col & 511;
		const uint32_t t = tiled_index(x, y);
		if (part == 0)
			solidXY[col] = solidT[t];
		reinterpret_cast<uint4 *>(colorXYZ + (size_t)col * 64)[part] = reinterpret_cast<const uint4 *>(colorT + (size_t)t * 64)[part];
	}

	// ---- Save ----
	// GameMap::IsSurface for all 64 voxels of a column at once (:65-83); the map edge has no neighbour (no wrap)
	__device__ __forceinline__ uint64_t surface_mask(const uint64_t *__restrict__ solidT, int x, int y) {
		const uint64_t S = solidT[tiled_index(x, y)];
		uint64_t open = 1ull;                  // z == 0
		open |= ~(S << 1);                     // air above (z - 1)
		open |= ~(S >> 1) & ~(1ull << 63);     // air below (z + 1), not for z = 63
		if (x > 0)
			open |= ~solidT[tiled_index(x - 1, y)];
		if (x < OSB200_MAP_W - 1)
			open |= ~solidT[tiled_index(x + 1, y)];
		if (y > 0)
			open |= ~solidT[tiled_index(x, y - 1)];
		if (y < OSB200_MAP_H - 1)
			open |= ~solidT[tiled_index(x, y + 1)];
		return S & open;
	}

	// The span loop of GameMap::Save (:105-155) for one column; kWrite = false only counts bytes.
	template <bool kWrite>
	__device__ __forceinline__ uint32_t save_column(uint64_t S, uint64_t F, const uint32_t *__restrict__ col, uint8_t *out) {
		const int d = 64;
		auto solid = [&](int z) { return (S >> z) & 1ull; };
		auto surf = [&](int z) { return (F >> z) & 1ull; };
		uint32_t n = 0;
		int k = 0;
		while (k < d) {
			const int airStart = k;
			while (k < d && !solid(k))
				++k;
			const int topStart = k;
			while (k < d && surf(k))
				++k;
			const int topEnd = k;
			while (k < d && solid(k) && !surf(k))
				++k;
			const int bottomStart = k;
			int z = k;
			while (z < d && surf(z))
				++z;
			if (z != d)
				while (surf(k)) // stops: z found a non-surface voxel below d
					++k;
			const int bottomEnd = k;
			const int topLen = topEnd - topStart, bottomLen = bottomEnd - bottomStart;
			const int colors = topLen + bottomLen;
			if (kWrite) {
				out[n + 0] = (uint8_t)(k == d ? 0 : colors + 1);
				out[n + 1] = (uint8_t)topStart;
				out[n + 2] = (uint8_t)(topEnd - 1);
				out[n + 3] = (uint8_t)airStart;
			}
			n += 4;
			for (int pass = 0; pass < 2; pass++) {
				const int from = pass ? bottomStart : topStart, len = pass ? bottomLen : topLen;
				for (int i = 0; i < len; i++) {
					if (kWrite) {
						const uint32_t c = col[(from + i) & 63]; // WriteColor (:85-90): R, G, B, then the health byte
						out[n + 0] = (uint8_t)(c >> 16);
						out[n + 1] = (uint8_t)(c >> 8);
						out[n + 2] = (uint8_t)c;
						out[n + 3] = (uint8_t)(c >> 24);
					}
					n += 4;
				}
			}
		}
		return n;
	}

	__global__ void k_vxl_measure(const uint64_t *__restrict__ solidT, uint32_t *__restrict__ sizes) {
		const int k = blockIdx.x * blockDim.x + threadIdx.x;
		if (k >= kCols)
			return;
		const int x = k & 511, y = k >> 9; // file order: y outer, x inner
		sizes[k] = save_column<false>(solidT[tiled_index(x, y)], surface_mask(solidT, x, y), nullptr, nullptr);
	}

	// exclusive scan of kCols values: 256 blocks x 1024 values (local scan + block sums), then the offsets are added
	__global__ void __launch_bounds__(256) k_scan_local(uint32_t *__restrict__ v, uint32_t *__restrict__ blockSums) {
		__shared__ uint32_t s[256];
		const int base = blockIdx.x * 1024 + threadIdx.x * 4;
		uint32_t a[4], sum = 0;
#pragma unroll
		for (int i = 0; i < 4; i++) {
			a[i] = v[base + i];
			sum += a[i];
		}
		s[threadIdx.x] = sum;
		__syncthreads();
		for (int off = 1; off < 256; off <<= 1) {
			const uint32_t add = threadIdx.x >= off ? s[threadIdx.x - off] : 0;
			__syncthreads();
			s[threadIdx.x] += add;
			__syncthreads();
		}
		uint32_t run = s[threadIdx.x] - sum;
#pragma unroll
		for (int i = 0; i < 4; i++) {
			const uint32_t t = a[i];
			v[base + i] = run;
			run += t;
		}
		if (threadIdx.x == 255)
			blockSums[blockIdx.x] = s[255];
	}
	__global__ void __launch_bounds__(256) k_scan_blocks(uint32_t *__restrict__ blockSums, uint32_t *__restrict__ total) {
		__shared__ uint32_t s[256];
		const uint32_t mine = blockSums[threadIdx.x];
		s[threadIdx.x] = mine;
		__syncthreads();
		for (int off = 1; off < 256; off <<= 1) {
			const uint32_t add = threadIdx.x >= off ? s[threadIdx.x - off] : 0;
			__syncthreads();
			s[threadIdx.x] += add;
			__syncthreads();
		}
		blockSums[threadIdx.x] = s[threadIdx.x] - mine;
		if (threadIdx.x == 255)
			*total = s[255];
	}

	__global__ void k_vxl_emit(const uint64_t *__restrict__ solidT, const uint32_t *__restrict__ colorT,
	                           const uint32_t *__restrict__ offsets, const uint32_t *__restrict__ blockSums,
	                           uint8_t *__restrict__ out, uint32_t cap) {
		const int k = blockIdx.x * blockDim.x + threadIdx.x;
		if (k >= kCols)
			return;
		const int x = k & 511, y = k >> 9;
		const uint32_t t = tiled_index(x, y);
		const uint32_t off = offsets[k] + blockSums[k >> 10];
		const uint64_t S = solidT[t], F = surface_mask(solidT, x, y);
		if (off + save_column<false>(S, F, nullptr, nullptr) > cap)
			return; // the caller's buffer is too small: it gets the size back and nothing past cap is written
		save_column<true>(S, F, colorT + (size_t)t * 64, out + off);
	}

	static int vxl_init(osb200_ctx *c) {
		if (c->vxl)
			return 0;
		auto *V = new VxlState();
		c->vxl = V;
		OSB_CUDA(cudaMalloc(&V->d_start, (kCols + 1) * sizeof(uint32_t)));
		OSB_CUDA(cudaMalloc(&V->d_sizes, (kCols + 1) * sizeof(uint32_t)));
		OSB_CUDA(cudaMalloc(&V->d_blockSums, 257 * sizeof(uint32_t)));
		OSB_CUDA(cudaMalloc(&V->d_err, sizeof(int)));
		return 0;
	}
	void vxl_destroy(osb200_ctx *c) {
		VxlState *V = c->vxl;
		if (!V)
			return;
		cudaFree(V->d_file);
		cudaFree(V->d_jumpA);
		cudaFree(V->d_jumpB);
		cudaFree(V->d_start);
		cudaFree(V->d_sizes);
		cudaFree(V->d_blockSums);
		cudaFree(V->d_err);
		delete V;
		c->vxl = nullptr;
	}
	void map_replaced_hooks(osb200_ctx *c); // api.cu
} // namespace osb

extern "C" {

int osb200_map_load_vxl(osb200_ctx *c, const uint8_t *data, size_t len, uint64_t *host_solid, uint32_t *host_color) {
	using namespace osb;
	if (!c || !data) {
		set_error("osb200_map_load_vxl: null argument");
		return -1;
	}
	if (len < 4 || len > (1ull << 31)) {
		set_error("osb200_map_load_vxl: truncated VXL stream");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (vxl_init(c))
		return -2;
	VxlState *V = c->vxl;
	const uint32_t nWords = (uint32_t)(len / 4); // every record is a multiple of 4 bytes; a ragged tail is never read
	if (len > V->fileCap) {
		cudaFree(V->d_file);
		V->d_file = nullptr;
		V->fileCap = 0;
		OSB_CUDA(cudaMalloc(&V->d_file, len + (len >> 2) + 2048));
		V->fileCap = len + (len >> 2);
	}
	if ((size_t)nWords + 1 > V->jumpCap) {
		cudaFree(V->d_jumpA);
		cudaFree(V->d_jumpB);
		V->d_jumpA = V->d_jumpB = nullptr;
		V->jumpCap = 0;
		const size_t n = (size_t)nWords + (nWords >> 2) + 64;
		OSB_CUDA(cudaMalloc(&V->d_jumpA, n * sizeof(uint32_t)));
		OSB_CUDA(cudaMalloc(&V->d_jumpB, n * sizeof(uint32_t)));
		V->jumpCap = n;
	}
	cudaStream_t st = c->stream;
	OSB_CUDA(cudaMemcpyAsync(V->d_file, data, len, cudaMemcpyHostToDevice, st));
	OSB_CUDA(cudaMemsetAsync(V->d_err, 0, sizeof(int), st));
	// every column start is "unknown" (kBad) until a doubling round writes it: a short stream leaves most of them that
	// way (the rounds only run nWords + 1 threads), and the decoder must see kBad there, not stale or uninitialised words
	OSB_CUDA(cudaMemsetAsync(V->d_start, 0xff, (kCols + 1) * sizeof(uint32_t), st));
	OSB_CUDA(cudaMemsetAsync(V->d_start, 0, sizeof(uint32_t), st)); // column 0 starts at word 0
	OSB_CUDA(cudaMemsetAsync(c->d_color, 0, (size_t)kCols * 64 * 4, st));
	const unsigned blocks = (nWords + 1 + 255) / 256;
	{
		KLaunch kl(c, K_VXL_INDEX);
		k_vxl_jump<<<blocks, 256, 0, st>>>(V->d_file, nWords, V->d_jumpA);
		uint32_t *in = V->d_jumpA, *out = V->d_jumpB;
		for (uint32_t have = 1; have < (uint32_t)kCols + 1; have <<= 1) { // 19 rounds: kCols + 1 starts (the last = end)
			k_vxl_double<<<blocks, 256, 0, st>>>(in, out, nWords, V->d_start, have, (uint32_t)kCols + 1);
			std::swap(in, out);
			c->launches++;
		}
	}
	{
		KLaunch kl(c, K_VXL_DECODE);
		k_vxl_decode<<<kCols / 128, 128, 0, st>>>(V->d_file, nWords, V->d_start, c->d_solid, c->d_color, V->d_err);
	}
	OSB_CUDA(cudaGetLastError());
	int err = 0;
	OSB_CUDA(cudaMemcpyAsync(&err, V->d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
	OSB_CUDA(cudaStreamSynchronize(st));
	if (err) {
		c->mapLoaded = false;
		set_error("osb200_map_load_vxl: truncated or malformed VXL stream");
		return -1;
	}
	if (host_solid || host_color) {
		const size_t sb = (size_t)kCols * 8, cb = (size_t)kCols * 64 * 4;
		if (c->ensure_scratch(sb + cb))
			return -2;
		auto *dS = reinterpret_cast<uint64_t *>(c->d_scratch);
		auto *dC = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(c->d_scratch) + sb);
		{
			KLaunch kl(c, K_MAP_REPACK);
			k_map_untile<<<kCols / 16, 256, 0, st>>>(c->d_solid, c->d_color, dS, dC);
		}
		if (host_solid)
			OSB_CUDA(cudaMemcpyAsync(host_solid, dS, sb, cudaMemcpyDeviceToHost, st));
		if (host_color)
			OSB_CUDA(cudaMemcpyAsync(host_color, dC, cb, cudaMemcpyDeviceToHost, st));
		OSB_CUDA(cudaStreamSynchronize(st));
	}
	map_replaced_hooks(c);
	return 0;
}

long long osb200_map_save_vxl(osb200_ctx *c, uint8_t *out, size_t cap) {
	using namespace osb;
	if (!c || (!out && cap)) {
		set_error("osb200_map_save_vxl: null argument");
		return -1;
	}
	if (!c->mapLoaded) {
		set_error("osb200_map_save_vxl: no map uploaded");
		return -1;
	}
	if (c->anyDirty) {
		set_error("osb200_map_save_vxl: unflushed edits (call osb200_map_flush first)");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (vxl_init(c))
		return -2;
	VxlState *V = c->vxl;
	cudaStream_t st = c->stream;
	if (cap > (1ull << 31))
		cap = 1ull << 31;
	uint8_t *d_out = nullptr;
	if (cap) {
		if (c->ensure_scratch(cap))
			return -2;
		d_out = reinterpret_cast<uint8_t *>(c->d_scratch);
	}
	{
		KLaunch kl(c, K_VXL_ENCODE);
		k_vxl_measure<<<kCols / 128, 128, 0, st>>>(c->d_solid, V->d_sizes);
		k_scan_local<<<256, 256, 0, st>>>(V->d_sizes, V->d_blockSums);
		k_scan_blocks<<<1, 256, 0, st>>>(V->d_blockSums, V->d_blockSums + 256);
		if (cap)
			k_vxl_emit<<<kCols / 128, 128, 0, st>>>(c->d_solid, c->d_color, V->d_sizes, V->d_blockSums, d_out, (uint32_t)cap);
		c->launches += 3;
	}
	OSB_CUDA(cudaGetLastError());
	uint32_t total = 0;
	OSB_CUDA(cudaMemcpyAsync(&total, V->d_blockSums + 256, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	OSB_CUDA(cudaStreamSynchronize(st));
	if (cap) {
		OSB_CUDA(cudaMemcpyAsync(out, d_out, total < cap ? total : cap, cudaMemcpyDeviceToHost, st));
		OSB_CUDA(cudaStreamSynchronize(st));
	}
	return (long long)total;
}

} // extern "C"
