// Multi-GPU below the language bindings: one osb200_ctx per device of ONE process, a NCCL communicator over them, and
// the three things the path needs across GPUs:
//   * the map replicated on every device: uploaded over PCIe ONCE (device 0) and copied to the peers with ncclBroadcast
//     over NVLink; edits the same way -- the dirty columns are staged and uploaded once, their records broadcast, and
//     every device scatters them into its resident map (osb200_group_map_flush);
//   * independent units sharded by contiguous range with NO data-path collective: views (osb200_group_render) and rays
//     (osb200_group_castray2), one host thread per device so that the per-view host set-up runs in parallel too;
//   * the optional gather of the finished frames to one device (osb200_group_render_gather): grouped ncclSend / ncclRecv,
//     only the root receives.
// The reference has nothing distributed.  NCCL is bound at run time (dlopen "libnccl.so.2"): a
// single-GPU user of libosb200.so never loads it, and a process that already carries a NCCL (PyTorch) shares that copy.
#include "ctx.h"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <thread>

namespace osb {
	void map_mark_resident(osb200_ctx *c);
	size_t dirty_record_bytes();
	int map_scatter_staged(osb200_ctx *c, int n);
	int render_device_frames(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
	                         float fogDist, uint32_t **d_fb, float **d_depth); // terrain.cu
} // namespace osb

namespace {
	struct Nccl {
		void *lib = nullptr;
		ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
		ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
		ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
		ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
		ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
		ncclResult_t (*GroupStart)() = nullptr;
		ncclResult_t (*GroupEnd)() = nullptr;
		const char *(*GetErrorString)(ncclResult_t) = nullptr;
		bool load() {
			if (lib)
				return true;
			lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
			if (!lib)
				lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
			if (!lib)
				return false;
#define SYM(field, name) field = reinterpret_cast<decltype(field)>(dlsym(lib, name))
			SYM(CommInitAll, "ncclCommInitAll");
			SYM(CommDestroy, "ncclCommDestroy");
			SYM(Broadcast, "ncclBroadcast");
			SYM(Send, "ncclSend");
			SYM(Recv, "ncclRecv");
			SYM(GroupStart, "ncclGroupStart");
			SYM(GroupEnd, "ncclGroupEnd");
			SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
			return CommInitAll && CommDestroy && Broadcast && Send && Recv && GroupStart && GroupEnd && GetErrorString;
		}
	};
	Nccl g_nccl;
} // namespace

struct osb200_group {
	std::vector<osb200_ctx *> ctx;
	std::vector<ncclComm_t> comm;
	std::string err; // first error of a per-device worker thread
};

#define OSB_NCCL(expr)                                                                                     \
	do {                                                                                                   \
		ncclResult_t r_ = (expr);                                                                          \
		if (r_ != ncclSuccess) {                                                                           \
			osb::set_error("%s failed: %s (%s:%d)", #expr, g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
			return -2;                                                                                     \
		}                                                                                                  \
	} while (0)

namespace {
	// one host thread per device; the first failing worker's message becomes the call's error (set_error is thread-local)
	template <class F> int for_each_device(osb200_group *g, F fn) {
		const int G = (int)g->ctx.size();
		std::vector<int> rc(G, 0);
		std::vector<std::string> msg(G);
		std::vector<std::thread> th;
		for (int i = 0; i < G; i++)
			th.emplace_back([&, i] {
				rc[i] = fn(i);
				if (rc[i] < 0)
					msg[i] = osb200_last_error();
			});
		for (auto &t : th)
			t.join();
		for (int i = 0; i < G; i++)
			if (rc[i] < 0) {
				osb::set_error("device %d: %s", g->ctx[i]->device, msg[i].c_str());
				return rc[i];
			}
		return 0;
	}
	inline void shard(size_t n, int i, int G, size_t &lo, size_t &hi) {
		lo = n * (size_t)i / (size_t)G;
		hi = n * (size_t)(i + 1) / (size_t)G;
	}
} // namespace

extern "C" {

int osb200_group_create(const osb200_opts *opts, const int *devices, int nDevices, osb200_group **out) {
	if (!out || !devices || nDevices < 1) {
		osb::set_error("osb200_group_create: bad argument");
		return -1;
	}
	*out = nullptr;
	auto *g = new osb200_group();
	for (int i = 0; i < nDevices; i++) {
		osb200_opts o{};
		if (opts)
			o = *opts;
		o.device = devices[i];
		osb200_ctx *c = nullptr;
		const int rc = osb200_create(&o, &c);
		if (rc) {
			for (osb200_ctx *p : g->ctx)
				osb200_destroy(p);
			delete g;
			return rc;
		}
		g->ctx.push_back(c);
	}
	if (nDevices > 1) {
		if (!g_nccl.load()) {
			osb::set_error("osb200_group_create: libnccl.so.2 not found (%s)", dlerror());
			osb200_group_destroy(g);
			return -3;
		}
		g->comm.resize(nDevices);
		ncclResult_t r = g_nccl.CommInitAll(g->comm.data(), nDevices, devices);
		if (r != ncclSuccess) {
			osb::set_error("ncclCommInitAll failed: %s", g_nccl.GetErrorString(r));
			g->comm.clear();
			osb200_group_destroy(g);
			return -2;
		}
	}
	*out = g;
	return 0;
}

void osb200_group_destroy(osb200_group *g) {
	if (!g)
		return;
	for (size_t i = 0; i < g->comm.size(); i++) {
		cudaSetDevice(g->ctx[i]->device);
		cudaStreamSynchronize(g->ctx[i]->stream);
		g_nccl.CommDestroy(g->comm[i]);
	}
	for (osb200_ctx *c : g->ctx)
		osb200_destroy(c);
	delete g;
}

int osb200_group_size(const osb200_group *g) { return g ? (int)g->ctx.size() : 0; }
osb200_ctx *osb200_group_ctx(osb200_group *g, int i) { return g && i >= 0 && i < (int)g->ctx.size() ? g->ctx[i] : nullptr; }

int osb200_group_map_upload(osb200_group *g, const uint64_t *solid, const uint32_t *color) {
	if (!g) {
		osb::set_error("osb200_group_map_upload: no group");
		return -1;
	}
	// PCIe once ...
	if (int rc = osb200_map_upload(g->ctx[0], solid, color))
		return rc;
	if (g->ctx.size() == 1)
		return 0;
	// ... then NVLink: the tiled arrays of device 0 go to every peer (the two arrays are one allocation on every device)
	const size_t ncol = (size_t)OSB200_MAP_W * OSB200_MAP_H;
	return for_each_device(g, [&](int i) -> int {
		osb200_ctx *c = g->ctx[i];
		OSB_CUDA(cudaSetDevice(c->device));
		OSB_NCCL(g_nccl.GroupStart());
		OSB_NCCL(g_nccl.Broadcast(c->d_solid, c->d_solid, ncol * 8, ncclUint8, 0, g->comm[i], c->stream));
		OSB_NCCL(g_nccl.Broadcast(c->d_color, c->d_color, ncol * 64 * 4, ncclUint8, 0, g->comm[i], c->stream));
		OSB_NCCL(g_nccl.GroupEnd());
		OSB_CUDA(cudaStreamSynchronize(c->stream));
		if (i)
			osb::map_mark_resident(c);
		return 0;
	});
}

int osb200_group_map_mark_dirty(osb200_group *g, int x, int y, int z) {
	if (!g)
		return -1;
	for (osb200_ctx *c : g->ctx) // every replica tracks the edit (its shadow / radiosity / ambient marks are per device)
		if (int rc = osb200_map_mark_dirty(c, x, y, z))
			return rc;
	return 0;
}

int osb200_group_map_flush(osb200_group *g, const uint64_t *solid, const uint32_t *color) {
	if (!g) {
		osb::set_error("osb200_group_map_flush: no group");
		return -1;
	}
	// device 0: stage the dirty columns, one H2D, scatter (the ordinary single-device flush)
	const int n = osb200_map_flush(g->ctx[0], solid, color);
	if (n <= 0 || g->ctx.size() == 1)
		return n;
	// the records now sit in device 0's staging buffer: broadcast them, every peer scatters them into its own map
	const size_t bytes = (size_t)n * osb::dirty_record_bytes();
	int rc = for_each_device(g, [&](int i) -> int {
		osb200_ctx *c = g->ctx[i];
		OSB_CUDA(cudaSetDevice(c->device));
		if (i && c->ensure_stage(bytes))
			return -2;
		OSB_NCCL(g_nccl.Broadcast(c->d_stage, c->d_stage, bytes, ncclUint8, 0, g->comm[i], c->stream));
		if (i)
			return osb::map_scatter_staged(c, n);
		return 0;
	});
	return rc ? rc : n;
}

int osb200_group_render(osb200_group *g, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
                        float fogDist, uint32_t *host_bgrx, float *host_depth) {
	if (!g || nViews < 0) {
		osb::set_error("osb200_group_render: bad argument");
		return -1;
	}
	const int G = (int)g->ctx.size();
	return for_each_device(g, [&](int i) -> int {
		size_t lo, hi;
		shard((size_t)nViews, i, G, lo, hi);
		if (hi == lo)
			return 0;
		const size_t off = lo * (size_t)w * h;
		return osb200_render(g->ctx[i], scenes + lo, (int)(hi - lo), w, h, fogColor, fogDist, host_bgrx ? host_bgrx + off : nullptr,
		                     host_depth ? host_depth + off : nullptr);
	});
}

int osb200_group_render_gather(osb200_group *g, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
                               float fogDist, int root, uint32_t *d_bgrx_root, float *d_depth_root) {
	if (!g || nViews < 0 || root < 0 || root >= (int)g->ctx.size() || !d_bgrx_root) {
		osb::set_error("osb200_group_render_gather: bad argument");
		return -1;
	}
	const int G = (int)g->ctx.size();
	const size_t px = (size_t)w * h;
	return for_each_device(g, [&](int i) -> int {
		osb200_ctx *c = g->ctx[i];
		size_t lo, hi;
		shard((size_t)nViews, i, G, lo, hi);
		uint32_t *fb = nullptr;
		float *dp = nullptr;
		if (hi > lo)
			if (int rc = osb::render_device_frames(c, scenes + lo, (int)(hi - lo), w, h, fogColor, fogDist, &fb, &dp))
				return rc;
		OSB_CUDA(cudaSetDevice(c->device));
		if (G > 1)
			OSB_NCCL(g_nccl.GroupStart());
		if (i == root) {
			for (int p = 0; p < G; p++) {
				size_t plo, phi;
				shard((size_t)nViews, p, G, plo, phi);
				if (phi == plo)
					continue;
				if (p == root) {
					OSB_CUDA(cudaMemcpyAsync(d_bgrx_root + plo * px, fb, (phi - plo) * px * 4, cudaMemcpyDeviceToDevice, c->stream));
					if (d_depth_root)
						OSB_CUDA(cudaMemcpyAsync(d_depth_root + plo * px, dp, (phi - plo) * px * 4, cudaMemcpyDeviceToDevice, c->stream));
				} else {
					OSB_NCCL(g_nccl.Recv(d_bgrx_root + plo * px, (phi - plo) * px, ncclUint32, p, g->comm[i], c->stream));
					if (d_depth_root)
						OSB_NCCL(g_nccl.Recv(d_depth_root + plo * px, (phi - plo) * px, ncclFloat32, p, g->comm[i], c->stream));
				}
			}
		} else if (hi > lo) {
			OSB_NCCL(g_nccl.Send(fb, (hi - lo) * px, ncclUint32, root, g->comm[i], c->stream));
			if (d_depth_root)
				OSB_NCCL(g_nccl.Send(dp, (hi - lo) * px, ncclFloat32, root, g->comm[i], c->stream));
		}
		if (G > 1)
			OSB_NCCL(g_nccl.GroupEnd());
		OSB_CUDA(cudaStreamSynchronize(c->stream));
		return 0;
	});
}

int osb200_group_castray2(osb200_group *g, const float *v0, const float *dir, int maxSteps, osb200_rayhit *out, size_t n) {
	if (!g) {
		osb::set_error("osb200_group_castray2: no group");
		return -1;
	}
	const int G = (int)g->ctx.size();
	return for_each_device(g, [&](int i) -> int {
		size_t lo, hi;
		shard(n, i, G, lo, hi);
		if (hi == lo)
			return 0;
		return osb200_castray2(g->ctx[i], v0 + lo * 3, dir + lo * 3, maxSteps, out + lo, hi - lo);
	});
}

} // extern "C"
