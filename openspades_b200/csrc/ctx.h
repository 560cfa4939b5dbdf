// Internal state behind the opaque osb200_ctx handle (include/osb200.h).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/osb200.h"

namespace osb {

	void set_error(const char *fmt, ...);

#define OSB_CUDA(expr)                                                                             \
	do {                                                                                           \
		cudaError_t e_ = (expr);                                                                   \
		if (e_ != cudaSuccess) {                                                                   \
			osb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__,       \
			               __LINE__);                                                              \
			return -2;                                                                             \
		}                                                                                          \
	} while (0)

	// The solid bitmap is stored in 4x4-column tiles of 128 B (one L1/L2 line) so that a ray or a
	// DDA line walking in any horizontal direction stays ~4 steps inside one line, instead of the
	// reference's x-major [x][y] order (GameMap.h:160) where every x-step is a 4 KiB jump.
	__host__ __device__ inline uint32_t tiled_index(int x, int y) {
		return ((uint32_t)(x >> 2) * 128u + (uint32_t)(y >> 2)) * 16u + (uint32_t)(x & 3) * 4u +
		       (uint32_t)(y & 3);
	}

	// kernel classes reported by osb200_timing_read (index = osb200_kernel_time slot)
	enum KernelId {
		K_MAP_REPACK = 0,
		K_MAP_SCATTER,
		K_VIEW_PREP,
		K_WALK_LINES,
		K_GATHER_FOG,
		K_CASTRAY2,
		K_SHADOW_BITMAP,
		K_RADIOSITY,
		K_RAD_SELECT,
		K_CASTRAY,
		K_AMBIENT,
		K_SCENE_FINISH,
		K_MODEL_SETUP,
		K_MODEL_SPLAT,
		K_LINE_SETUP,
		K_LIGHTS,
		K_VXL_INDEX,
		K_VXL_DECODE,
		K_VXL_ENCODE,
		K_LINKS_FLOOD,
		K_MINIMAP,
		K_CLUSTER_UNION,
		K_VIEW_PIPELINE,
		K_COUNT
	};
	const char *kernel_name(int id);

	// osb200_castray2's chunk pipeline (castray2.cu)
	constexpr int kPipeSlots = 3;
	constexpr size_t kPipeRays = (size_t)1 << 20; // rays per chunk: 64 MB per slot
	constexpr size_t kStagedRays = (size_t)1 << 18; // rays per chunk when the caller's buffers are pageable: 16 MB per slot
	struct RayPipe {
		cudaStream_t in = nullptr, out = nullptr;
		cudaEvent_t uploaded[kPipeSlots] = {}, computed[kPipeSlots] = {}, downloaded[kPipeSlots] = {};
		void *h_ring = nullptr; // pinned staging ring of the pageable path (kPipeSlots slots of kStagedRays rays)
		size_t h_ring_bytes = 0;
	};
	struct TerrainState;   // terrain.cu
	struct RadiosityState; // radiosity.cu
	struct AmbientState;   // ambient.cu
	struct ModelState;     // models.cu
	struct VxlState;       // vxl.cu
	struct LinksState;     // links.cu
} // namespace osb

struct osb200_ctx;
namespace osb {
	// terrain.cu, shared with models.cu (the full EndScene: terrain -> models -> lights -> fog)
	// castray2.cu: the solid map in [x][y] order (index (x & 511) << 9 | (y & 511)), up to date with the resident map
	int ray_map(osb200_ctx *c, const uint64_t **out);
	int render_impl(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
	                float fogDist, uint32_t *d_fb, float *d_depth, uint32_t *host_fb, float *host_depth, bool deferFog);
	int terrain_out_buffers(osb200_ctx *c, size_t n, uint32_t **fb, float **depth);

	// one kept dynamic light of a scene (models.cu fills it: AddLight's culls and rectangle, SWRenderer.cpp:578-649)
	struct LightDev {
		int minX, minY, maxX, maxY;
		float center[3]; // light origin in view space
		int r, g, b;     // ToFixedFactor8
		float invRadius2;
		uint32_t tabBase; // vx[fw] then vy[fh] of this light inside the light table buffer
	};
	// The tail of EndScene for a frame with models and / or lights in ONE pass over the frame (terrain.cu): winner of
	// the splat keys (keys may be null), the lights in submission order (nLights may be 0), then the fog.
	int terrain_finish_scene(osb200_ctx *c, int w, int h, const float fogColor[3], uint32_t *d_fb, float *d_depth,
	                         const unsigned long long *d_keys, const uint32_t *d_splatColor, const LightDev *d_lights,
	                         int nLights, const float *d_lightTabs);
	// order-preserving map float -> uint32 (and back), so that one unsigned atomicMin implements `zval < *db`
	__device__ __forceinline__ uint32_t depth_key(float f) {
		const uint32_t u = __float_as_uint(f);
		return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
	}
	__device__ __forceinline__ float key_depth(uint32_t k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); }

	// RAII bracket around one kernel launch: counts it and, when timing is on, records CUDA events
	// on the context stream before and after it.
	struct KLaunch {
		osb200_ctx *c;
		int slot;
		cudaStream_t st;
		KLaunch(osb200_ctx *c, int kid, cudaStream_t stream = nullptr); // default: the context stream
		~KLaunch();
	};
} // namespace osb

struct osb200_ctx {
	osb200_opts opts{};
	int device = 0;
	cudaStream_t stream = nullptr;
	uint64_t launches = 0;
	int smCount = 148;

	// ---- per-kernel CUDA-event timing (osb200_timing_enable / osb200_timing_read) ----
	struct TimedLaunch {
		int kid;
		cudaEvent_t a, b;
	};
	bool timing = false;
	std::vector<TimedLaunch> timed;      // recorded, not yet read
	std::vector<cudaEvent_t> eventPool;  // recycled events
	cudaEvent_t take_event();

	// ---- resident map ----
	bool mapLoaded = false;
	uint64_t *d_solid = nullptr; // [tiled_index(x,y)] 2 MiB
	uint32_t *d_color = nullptr; // [(tiled_index(x,y)) * 64 + z] 64 MiB, 0xHHBBGGRR as in GameMap (same allocation)
	size_t l2PersistBytes = 0;   // L2 carve-out pinned to the map window (0 = not available)
	// dirty tracking (same 32-column words as SWFlatMapRenderer::updateMap, SWFlatMapRenderer.cpp:107-111)
	std::vector<uint32_t> dirtyWords; // 512*512/32
	bool anyDirty = false;
	// pinned staging for dirty-column re-upload
	void *h_stage = nullptr;
	size_t h_stage_bytes = 0;
	void *d_stage = nullptr;
	size_t d_stage_bytes = 0;

	// ---- SSE emulation tables (device copies; null in IEEE mode) ----
	uint32_t *d_rcp = nullptr;
	int rcpBits = 0;
	uint32_t *d_rsqrt = nullptr;
	int rsqrtBits = 0;

	unsigned long long *d_rayCursor = nullptr; // batch-wide work cursor of k_castray2
	// the solid map once more in GameMap's own [x][y] order, for the ray kernel (its address is three instructions instead
	// of eight for the tiled one, and the kernel is bound by exactly that kind of instruction); rebuilt when mapVersion moved
	uint64_t *d_solidXY = nullptr;
	uint64_t mapVersion = 0, solidXYVersion = ~0ull;
	osb::RayPipe *rayPipe = nullptr;           // streams / events of the host-buffer ray pipeline

	// ---- scratch for host-pointer convenience entry points ----
	void *d_scratch = nullptr;
	size_t d_scratch_bytes = 0;

	osb::TerrainState *terrain = nullptr;
	osb::RadiosityState *radiosity = nullptr;
	osb::AmbientState *ambient = nullptr;
	osb::ModelState *models = nullptr;
	osb::VxlState *vxl = nullptr;
	osb::LinksState *links = nullptr;

	int ensure_scratch(size_t bytes);
	int ensure_stage(size_t bytes);
};
