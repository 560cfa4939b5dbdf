// Context lifetime, error reporting and map residency for the osb200 C ABI (include/osb200.h).
//
// Reference behaviour being replaced:
//   SWRenderer::SetGameMap          Sources/Draw/SWRenderer.cpp:172-191
//   SWMapRenderer ctor / BuildRle   Sources/Draw/SWMapRenderer.cpp:107-203   (the six z-lists are pure bit
//       functions of a column's 64-bit mask and its neighbour's, so nothing is pre-built here: the
//       line kernel derives them from the resident bitmap -- see terrain.cu)
//   SWFlatMapRenderer dirty bitset  Sources/Draw/SWFlatMapRenderer.cpp:52-111
#include "ctx.h"

#include <algorithm>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

namespace osb {
	static thread_local std::string g_err;
	void set_error(const char *fmt, ...) {
		char buf[1024];
		va_list ap;
		va_start(ap, fmt);
		vsnprintf(buf, sizeof(buf), fmt, ap);
		va_end(ap);
		g_err = buf;
	}
	const char *kernel_name(int id) {
		static const char *names[K_COUNT] = {"k_map_repack", "k_map_scatter", "k_view_prep", "k_walk_lines",
		                                     "k_gather_fog", "k_castray2", "k_shadow_update", "k_radiosity",
		                                     "k_rad_select", "k_castray", "k_ao_chunk", "k_scene_finish", "k_model_setup", "k_model_splat",
		                                     "k_line_setup", "k_light_prep", "k_vxl_index", "k_vxl_decode", "k_vxl_encode", "k_links_flood", "k_minimap", "k_cluster_union", "k_view_pipeline"};
		return id >= 0 && id < K_COUNT ? names[id] : "?";
	}
	KLaunch::KLaunch(osb200_ctx *ctx, int kid, cudaStream_t stream) : c(ctx), slot(-1), st(stream ? stream : ctx->stream) {
		c->launches++;
		if (!c->timing)
			return;
		osb200_ctx::TimedLaunch t{kid, c->take_event(), c->take_event()};
		if (!t.a || !t.b)
			return;
		cudaEventRecord(t.a, st);
		c->timed.push_back(t);
		slot = (int)c->timed.size() - 1;
	}
	KLaunch::~KLaunch() {
		if (slot >= 0)
			cudaEventRecord(c->timed[slot].b, st);
	}
	void terrain_destroy(osb200_ctx *ctx);   // terrain.cu
	void radiosity_destroy(osb200_ctx *ctx); // radiosity.cu
	void radiosity_map_replaced(osb200_ctx *ctx);
	void radiosity_mark_dirty(osb200_ctx *ctx, int x, int y, int z);
	void ambient_destroy(osb200_ctx *ctx);   // ambient.cu
	void models_destroy(osb200_ctx *ctx);    // models.cu
	void vxl_destroy(osb200_ctx *ctx);       // vxl.cu
	void links_destroy(osb200_ctx *ctx);     // links.cu
	void links_map_replaced(osb200_ctx *ctx);
	void ambient_map_replaced(osb200_ctx *ctx);
	void ambient_mark_dirty(osb200_ctx *ctx, int x, int y, int z);
	// a whole new map is resident (osb200_map_upload / osb200_map_load_vxl): nothing is dirty, every derived volume is stale
	void map_replaced_hooks(osb200_ctx *c) {
		std::fill(c->dirtyWords.begin(), c->dirtyWords.end(), 0u);
		c->anyDirty = false;
		c->mapLoaded = true;
		c->mapVersion++;
		radiosity_map_replaced(c);
		ambient_map_replaced(c);
		links_map_replaced(c);
	}
} // namespace osb

cudaEvent_t osb200_ctx::take_event() {
	if (!eventPool.empty()) {
		cudaEvent_t e = eventPool.back();
		eventPool.pop_back();
		return e;
	}
	cudaEvent_t e = nullptr;
	if (cudaEventCreate(&e) != cudaSuccess)
		return nullptr;
	return e;
}

int osb200_ctx::ensure_scratch(size_t bytes) {
	if (bytes <= d_scratch_bytes)
		return 0;
	if (d_scratch)
		cudaFree(d_scratch);
	d_scratch = nullptr;
	d_scratch_bytes = 0;
	size_t want = bytes + (bytes >> 2);
	OSB_CUDA(cudaMalloc(&d_scratch, want));
	d_scratch_bytes = want;
	return 0;
}

int osb200_ctx::ensure_stage(size_t bytes) {
	if (bytes <= h_stage_bytes && bytes <= d_stage_bytes)
		return 0;
	size_t want = bytes + (bytes >> 1) + 4096;
	if (h_stage)
		cudaFreeHost(h_stage);
	if (d_stage)
		cudaFree(d_stage);
	h_stage = d_stage = nullptr;
	h_stage_bytes = d_stage_bytes = 0;
	OSB_CUDA(cudaMallocHost(&h_stage, want));
	h_stage_bytes = want;
	OSB_CUDA(cudaMalloc(&d_stage, want));
	d_stage_bytes = want;
	return 0;
}

// ---- kernels ----------------------------------------------------------------------------------

// [x][y] (reference order) -> 4x4-tiled order.  One thread per column; colour rows are 256 B so
// each thread moves 16 x uint4.
__global__ void k_map_repack(const uint64_t *__restrict__ solidXY, const uint32_t *__restrict__ colorXYZ,
                             uint64_t *__restrict__ solidT, uint32_t *__restrict__ colorT) {
	int col = blockIdx.x * (blockDim.x >> 4) + (threadIdx.x >> 4); // 16 threads per column
	int part = threadIdx.x & 15;
	if (col >= OSB200_MAP_W * OSB200_MAP_H)
		return;
	int x = col >> 9, y = col & 511;
	uint32_t t = osb::tiled_index(x, y);
	if (part == 0)
		solidT[t] = solidXY[col];
	const uint4 *src = reinterpret_cast<const uint4 *>(colorXYZ + (size_t)col * 64);
	uint4 *dst = reinterpret_cast<uint4 *>(colorT + (size_t)t * 64);
	dst[part] = src[part];
}

struct DirtyColumn { // 272 B staging record of one edited column
	uint32_t tiled;
	uint32_t pad;
	uint64_t solid;
	uint32_t color[64];
};

__global__ void k_map_scatter(const DirtyColumn *__restrict__ recs, int n, uint64_t *__restrict__ solidT,
                              uint32_t *__restrict__ colorT) {
	int i = blockIdx.x * (blockDim.x >> 4) + (threadIdx.x >> 4);
	int part = threadIdx.x & 15;
	if (i >= n)
		return;
	const DirtyColumn &r = recs[i];
	if (part == 0)
		solidT[r.tiled] = r.solid;
	const uint4 *src = reinterpret_cast<const uint4 *>(r.color);
	reinterpret_cast<uint4 *>(colorT + (size_t)r.tiled * 64)[part] = src[part];
}

__global__ void k_map_untile_solid(const uint64_t *__restrict__ solidT, uint64_t *__restrict__ solidXY) {
	int col = blockIdx.x * blockDim.x + threadIdx.x;
	if (col < OSB200_MAP_W * OSB200_MAP_H)
		solidXY[col] = solidT[osb::tiled_index(col >> 9, col & 511)];
}

namespace osb {
	// group.cu: the tiled arrays of `c` were filled by a broadcast from a peer device
	void map_mark_resident(osb200_ctx *c) { map_replaced_hooks(c); }
	size_t dirty_record_bytes() { return sizeof(DirtyColumn); }
	// group.cu: n dirty-column records arrived in c->d_stage by broadcast: scatter them, the host marks are settled
	int map_scatter_staged(osb200_ctx *c, int n) {
		if (n > 0) {
			KLaunch kl(c, K_MAP_SCATTER);
			k_map_scatter<<<(n + 15) / 16, 256, 0, c->stream>>>(reinterpret_cast<DirtyColumn *>(c->d_stage), n, c->d_solid,
			                                                    c->d_color);
			c->mapVersion++;
		}
		std::fill(c->dirtyWords.begin(), c->dirtyWords.end(), 0u);
		c->anyDirty = false;
		OSB_CUDA(cudaGetLastError());
		return 0;
	}
} // namespace osb

// ---- C ABI ------------------------------------------------------------------------------------

extern "C" {

const char *osb200_last_error(void) { return osb::g_err.c_str(); }

int osb200_create(const osb200_opts *opts, osb200_ctx **out) {
	if (!opts || !out) {
		osb::set_error("osb200_create: null argument");
		return -1;
	}
	*out = nullptr;
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0) {
		// no CPU fallback: the product path fails loudly without a device
		osb::set_error("osb200_create: no CUDA device available (%s); this library has no CPU fallback",
		               e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
		return -3;
	}
	if (opts->device < 0 || opts->device >= ndev) {
		osb::set_error("osb200_create: device %d out of range (0..%d)", opts->device, ndev - 1);
		return -1;
	}
	if (opts->numerics < OSB200_NUM_IEEE || opts->numerics > OSB200_NUM_SSE2) {
		osb::set_error("osb200_create: bad numerics mode %d", opts->numerics);
		return -1;
	}
	const uint32_t *rcpTab = opts->rcp_table, *rsqrtTab = opts->rsqrt_table;
	int rcpBits = opts->rcp_bits, rsqrtBits = opts->rsqrt_bits;
	if (opts->numerics != OSB200_NUM_IEEE) {
		if (!rcpTab || !rsqrtTab) {
			// measure this host's RCPSS/RSQRTSS (the CPU the reference SWRenderer would run on): an exhaustive scan of
			// ~10^8 evaluations, done ONCE per process and shared by every context (ranks, groups, tests)
			struct Cal {
				std::vector<uint32_t> rcp, rsqrt;
				int rcpBits = 0, rsqrtBits = 0, rc = 0;
				std::string err;
			};
			static Cal cal;
			static std::once_flag once;
			std::call_once(once, [] {
				cal.rcp.resize(1u << 16);
				cal.rsqrt.resize(2u << 16);
				cal.rc = osb200_calibrate_sse(cal.rcp.data(), &cal.rcpBits, cal.rsqrt.data(), &cal.rsqrtBits, 16);
				if (cal.rc)
					cal.err = osb200_last_error();
			});
			if (cal.rc) {
				osb::set_error("osb200_create: the host's RCPSS/RSQRTSS could not be modelled (%s); the SSE / SSE2 numerics "
				               "variants reproduce the reference as it runs on THIS CPU and need that model -- pass tables "
				               "measured elsewhere in osb200_opts, or use numerics = OSB200_NUM_IEEE (the reference built "
				               "without SSE), which needs none",
				               cal.err.c_str());
				return cal.rc;
			}
			rcpTab = cal.rcp.data();
			rsqrtTab = cal.rsqrt.data();
			rcpBits = cal.rcpBits;
			rsqrtBits = cal.rsqrtBits;
		}
		if (rcpBits < 1 || rcpBits > 16 || rsqrtBits < 1 || rsqrtBits > 16) {
			osb::set_error("osb200_create: bad rcpss/rsqrtss table sizes");
			return -1;
		}
	}
	auto *c = new osb200_ctx();
	c->opts = *opts;
	if (c->opts.sw_threads <= 0)
		c->opts.sw_threads = 4;
	if (c->opts.sw_threads > 32)
		c->opts.sw_threads = 32; // SWUtils.h:53
	if (c->opts.undersampling != 2 && c->opts.undersampling != 4)
		c->opts.undersampling = 1;
	if (c->opts.radiosity_mode != 2)
		c->opts.radiosity_mode = 1;
	c->device = opts->device;
	c->dirtyWords.assign(OSB200_MAP_W * OSB200_MAP_H / 32, 0);
	auto fail = [&](int rc) {
		osb200_destroy(c);
		return rc;
	};
#define TRY(expr)                                                                                  \
	do {                                                                                           \
		cudaError_t e_ = (expr);                                                                   \
		if (e_ != cudaSuccess) {                                                                   \
			osb::set_error("%s failed: %s", #expr, cudaGetErrorString(e_));                        \
			return fail(-2);                                                                       \
		}                                                                                          \
	} while (0)
	TRY(cudaSetDevice(c->device));
	cudaDeviceProp prop;
	TRY(cudaGetDeviceProperties(&prop, c->device));
	c->smCount = prop.multiProcessorCount;
	TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
	{
		// one allocation for bitmap (2 MiB) + colours (64 MiB) so that a single L2 access-policy window covers
		// the whole resident map: the DDA / ray loops do dependent random reads into it while GBs of strips,
		// rays and frames stream past, so it is pinned in L2 (persisting hits, streaming misses elsewhere)
		const size_t sb = sizeof(uint64_t) * OSB200_MAP_W * OSB200_MAP_H;
		const size_t cb = sizeof(uint32_t) * OSB200_MAP_W * OSB200_MAP_H * OSB200_MAP_D;
		void *base = nullptr;
		TRY(cudaMalloc(&base, sb + cb));
		c->d_solid = reinterpret_cast<uint64_t *>(base);
		c->d_color = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(base) + sb);
		TRY(cudaMalloc(&c->d_solidXY, sb)); // the ray kernel's copy of the solid map (castray2.cu); here, so that no ray call allocates
		const char *env = getenv("OSB200_L2_PERSIST");
		const bool want = env && atoi(env) != 0; // measured on B200: the carve-out costs more (strip write-combining capacity) than it saves -> opt-in only
		if (want && prop.persistingL2CacheMaxSize > 0 && prop.accessPolicyMaxWindowSize > 0) {
			const size_t win = std::min<size_t>(sb + cb, (size_t)prop.accessPolicyMaxWindowSize);
			const size_t carve = std::min<size_t>((size_t)prop.persistingL2CacheMaxSize, win);
			if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) == cudaSuccess) {
				cudaStreamAttrValue attr;
				memset(&attr, 0, sizeof(attr));
				attr.accessPolicyWindow.base_ptr = base;
				attr.accessPolicyWindow.num_bytes = win;
				attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)carve / (double)win);
				attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
				attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
				if (cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &attr) == cudaSuccess)
					c->l2PersistBytes = carve;
			}
			cudaGetLastError(); // the window is an optimisation: never fatal
		}
	}
	if (opts->numerics != OSB200_NUM_IEEE) {
		c->rcpBits = rcpBits;
		c->rsqrtBits = rsqrtBits;
		size_t nr = (size_t)1 << c->rcpBits, ns = (size_t)2 << c->rsqrtBits;
		TRY(cudaMalloc(&c->d_rcp, nr * 4));
		TRY(cudaMalloc(&c->d_rsqrt, ns * 4));
		TRY(cudaMemcpy(c->d_rcp, rcpTab, nr * 4, cudaMemcpyHostToDevice));
		TRY(cudaMemcpy(c->d_rsqrt, rsqrtTab, ns * 4, cudaMemcpyHostToDevice));
	}
	c->opts.rcp_table = c->opts.rsqrt_table = nullptr; // caller's pointers are not retained
#undef TRY
	*out = c;
	return 0;
}

void osb200_destroy(osb200_ctx *c) {
	if (!c)
		return;
	cudaSetDevice(c->device);
	if (c->stream)
		cudaStreamSynchronize(c->stream);
	osb::terrain_destroy(c);
	osb::radiosity_destroy(c);
	osb::ambient_destroy(c);
	osb::models_destroy(c);
	osb::vxl_destroy(c);
	osb::links_destroy(c);
	cudaFree(c->d_solid); // bitmap + colours are one allocation
	cudaFree(c->d_rcp);
	cudaFree(c->d_rsqrt);
	cudaFree(c->d_scratch);
	cudaFree(c->d_rayCursor);
	cudaFree(c->d_solidXY);
	if (c->rayPipe) {
		cudaStreamDestroy(c->rayPipe->in);
		cudaStreamDestroy(c->rayPipe->out);
		for (int i = 0; i < osb::kPipeSlots; i++) {
			cudaEventDestroy(c->rayPipe->uploaded[i]);
			cudaEventDestroy(c->rayPipe->computed[i]);
			cudaEventDestroy(c->rayPipe->downloaded[i]);
		}
		if (c->rayPipe->h_ring)
			cudaFreeHost(c->rayPipe->h_ring);
		delete c->rayPipe;
	}
	cudaFree(c->d_stage);
	if (c->h_stage)
		cudaFreeHost(c->h_stage);
	for (auto &t : c->timed) {
		cudaEventDestroy(t.a);
		cudaEventDestroy(t.b);
	}
	for (auto e : c->eventPool)
		cudaEventDestroy(e);
	if (c->stream)
		cudaStreamDestroy(c->stream);
	delete c;
}

int osb200_sync(osb200_ctx *c) {
	if (!c)
		return -1;
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

int osb200_timing_enable(osb200_ctx *c, int on) {
	if (!c)
		return -1;
	c->timing = on != 0;
	return 0;
}

int osb200_timing_read(osb200_ctx *c, osb200_kernel_time *out, int cap) {
	if (!c || (!out && cap > 0)) {
		osb::set_error("osb200_timing_read: null argument");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	osb200_kernel_time acc[osb::K_COUNT];
	for (int k = 0; k < osb::K_COUNT; k++) {
		acc[k].name = osb::kernel_name(k);
		acc[k].launches = 0;
		acc[k].total_ms = 0.0;
	}
	for (auto &t : c->timed) {
		float ms = 0.f;
		if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) {
			acc[t.kid].launches++;
			acc[t.kid].total_ms += ms;
		}
		c->eventPool.push_back(t.a);
		c->eventPool.push_back(t.b);
	}
	c->timed.clear();
	int n = 0;
	for (int k = 0; k < osb::K_COUNT && n < cap; k++)
		out[n++] = acc[k];
	return n;
}

void *osb200_stream(osb200_ctx *c) { return c ? (void *)c->stream : nullptr; }
uint64_t osb200_launch_count(osb200_ctx *c) { return c ? c->launches : 0; }

int osb200_map_upload(osb200_ctx *c, const uint64_t *solid, const uint32_t *color) {
	if (!c || !solid || !color) {
		osb::set_error("osb200_map_upload: null argument");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	const size_t ncol = (size_t)OSB200_MAP_W * OSB200_MAP_H;
	const size_t sb = ncol * 8, cb = ncol * 64 * 4;
	if (c->ensure_scratch(sb + cb))
		return -2;
	auto *dS = reinterpret_cast<uint64_t *>(c->d_scratch);
	auto *dC = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(c->d_scratch) + sb);
	OSB_CUDA(cudaMemcpyAsync(dS, solid, sb, cudaMemcpyHostToDevice, c->stream));
	OSB_CUDA(cudaMemcpyAsync(dC, color, cb, cudaMemcpyHostToDevice, c->stream));
	{
		osb::KLaunch kl(c, osb::K_MAP_REPACK);
		k_map_repack<<<(unsigned)(ncol / 16), 256, 0, c->stream>>>(dS, dC, c->d_solid, c->d_color);
	}
	OSB_CUDA(cudaGetLastError());
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	osb::map_replaced_hooks(c);
	return 0;
}

int osb200_map_clear(osb200_ctx *c) {
	if (!c)
		return -1;
	c->mapLoaded = false;
	std::fill(c->dirtyWords.begin(), c->dirtyWords.end(), 0u);
	c->anyDirty = false;
	return 0;
}

int osb200_map_mark_dirty(osb200_ctx *c, int x, int y, int z) {
	if (!c || x < 0 || x >= OSB200_MAP_W || y < 0 || y >= OSB200_MAP_H) {
		osb::set_error("osb200_map_mark_dirty: bad argument");
		return -1;
	}
	c->dirtyWords[(x + y * OSB200_MAP_W) >> 5] |= 1u << (x & 31);
	c->anyDirty = true;
	osb::radiosity_mark_dirty(c, x, y, z);
	osb::ambient_mark_dirty(c, x, y, z);
	return 0;
}

int osb200_map_flush(osb200_ctx *c, const uint64_t *solid, const uint32_t *color) {
	if (!c || !solid || !color) {
		osb::set_error("osb200_map_flush: null argument");
		return -1;
	}
	if (!c->mapLoaded) {
		osb::set_error("osb200_map_flush: no map uploaded");
		return -1;
	}
	if (!c->anyDirty)
		return 0;
	OSB_CUDA(cudaSetDevice(c->device));
	int n = 0;
	for (uint32_t w : c->dirtyWords)
		n += __builtin_popcount(w);
	if (c->ensure_stage((size_t)n * sizeof(DirtyColumn)))
		return -2;
	// the staging buffer may still be read by the previous flush's copy
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	auto *recs = reinterpret_cast<DirtyColumn *>(c->h_stage);
	int k = 0;
	for (size_t wi = 0; wi < c->dirtyWords.size(); wi++) {
		uint32_t w = c->dirtyWords[wi];
		while (w) {
			int b = __builtin_ctz(w);
			w &= w - 1;
			int lin = (int)wi * 32 + b; // x + y*512
			int x = lin & 511, y = lin >> 9;
			DirtyColumn &r = recs[k++];
			r.tiled = osb::tiled_index(x, y);
			r.pad = 0;
			r.solid = solid[(size_t)x * 512 + y];
			std::memcpy(r.color, color + ((size_t)x * 512 + y) * 64, 256);
		}
		c->dirtyWords[wi] = 0;
	}
	c->anyDirty = false;
	OSB_CUDA(cudaMemcpyAsync(c->d_stage, recs, (size_t)n * sizeof(DirtyColumn), cudaMemcpyHostToDevice,
	                         c->stream));
	{
		osb::KLaunch kl(c, osb::K_MAP_SCATTER);
		k_map_scatter<<<(n + 15) / 16, 256, 0, c->stream>>>(reinterpret_cast<DirtyColumn *>(c->d_stage), n,
		                                                    c->d_solid, c->d_color);
		c->mapVersion++;
	}
	OSB_CUDA(cudaGetLastError());
	return n;
}

int osb200_map_read_solid(osb200_ctx *c, uint64_t *solid) {
	if (!c || !solid || !c->mapLoaded) {
		osb::set_error("osb200_map_read_solid: bad argument or no map");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	const size_t ncol = (size_t)OSB200_MAP_W * OSB200_MAP_H;
	if (c->ensure_scratch(ncol * 8))
		return -2;
	k_map_untile_solid<<<(unsigned)(ncol / 256), 256, 0, c->stream>>>(
	  c->d_solid, reinterpret_cast<uint64_t *>(c->d_scratch));
	c->launches++;
	OSB_CUDA(cudaMemcpyAsync(solid, c->d_scratch, ncol * 8, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

} // extern "C"
