/*
 * osb200 -- C ABI of the B200-native voxel hot path for OpenSpades.
 *
 * The reference (yvt/openspades) has no FFI for this path: the software renderer is the C++
 * class draw::SWRenderer behind client::IRenderer (Sources/Client/IRenderer.h:98-174,
 * Sources/Draw/SWRenderer.h:48-219) and the ray query is the member GameMap::CastRay2
 * (Sources/Client/GameMap.h:150-157).  This header is the boundary a maintainer binds instead:
 * plain pointers and sizes, opaque context handle, int status (0 = ok, <0 = error, text via
 * osb200_last_error()).  Each entry point names the reference interface it replaces.
 * The host-side C++ class with SWRenderer's surface that sits on top of it is
 * openspades_b200/host/B200Renderer.h; INTEGRATION.md shows the reference-side wiring.
 *
 * Every function requires a CUDA device (sm_100a).  There is no CPU fallback.
 * A context is thread-compatible (one caller at a time), like SWRenderer itself.
 */
#ifndef OSB200_H
#define OSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OSB200_API __attribute__((visibility("default")))

#define OSB200_MAP_W 512 /* GameMap::DefaultWidth  (GameMap.h:45-49) */
#define OSB200_MAP_H 512
#define OSB200_MAP_D 64

typedef struct osb200_ctx osb200_ctx;

/* Which of the reference's numerically different builds of the SW path to reproduce
 * (Sources/Draw/SWFeatureLevel.h:25-121). */
enum osb200_numerics {
	OSB200_NUM_IEEE = 0, /* SWFeatureLevel::None built without SSE: fastRcp = 1/x, scalar shading, truncating fog */
	OSB200_NUM_SSE = 1,  /* SWFeatureLevel::SSE: rcpss/rsqrtss approximations, scalar shading, truncating fog */
	OSB200_NUM_SSE2 = 2  /* SWFeatureLevel::SSE2 (the reference's default): rcpss/rsqrtss, SSE2 shading, rounding fog */
};

typedef struct osb200_opts {
	int device;       /* CUDA device ordinal */
	int numerics;     /* enum osb200_numerics */
	int sw_threads;   /* emulated r_swNumThreads (SWRenderer.cpp:41): the reference's per-thread float
	                     recurrences in RenderFinal / ApplyFog make the frame depend on it. 0 -> 4 */
	int undersampling; /* r_swUndersampling (SWMapRenderer.cpp:37): 1, 2 or 4. 0 -> 1 */
	int radiosity_mode; /* r_radiosity (GLSettings.cpp:60): 1 = sqrt-compressed, 2 = linear. 0 -> 1 */
	/* rcpss / rsqrtss lookup tables for OSB200_NUM_SSE/SSE2: the result bits for every input whose
	 * mantissa's top `rcp_bits` (`rsqrt_bits`) bits are i and whose exponent is 0 (rcp) or 0/1 (rsqrt,
	 * 2 << rsqrt_bits entries, odd exponent second).  NULL -> osb200_create measures them itself on
	 * this host with osb200_calibrate_sse (i.e. matches the reference SWRenderer run on this CPU);
	 * pass tables to match another CPU.  Ignored in IEEE mode. */
	const uint32_t *rcp_table;
	int rcp_bits;
	const uint32_t *rsqrt_table;
	int rsqrt_bits;
} osb200_opts;

/* Fields of client::SceneDefinition the terrain path reads (Sources/Client/SceneDefinition.h:26-33). */
typedef struct osb200_scene {
	float origin[3];  /* viewOrigin */
	float axis[3][3]; /* viewAxis[0] = right, [1] = up, [2] = front */
	float fovX, fovY; /* radians */
	float zNear, zFar;
} osb200_scene;

/* Same fields, order and offsets as GameMap::RayCastResult (GameMap.h:150-156); 40 bytes. */
typedef struct osb200_rayhit {
	uint8_t hit, startSolid, pad_[2];
	float hitPos[3];
	int32_t hitBlock[3];
	int32_t normal[3];
} osb200_rayhit;

/* ---- lifetime / errors (replaces SWRenderer ctor/Init/Shutdown, SWRenderer.cpp:45-170) ---- */
OSB200_API const char *osb200_last_error(void);
OSB200_API int osb200_create(const osb200_opts *opts, osb200_ctx **out);
OSB200_API void osb200_destroy(osb200_ctx *ctx);
OSB200_API int osb200_sync(osb200_ctx *ctx);
/* Measures the host CPU's RCPSS/RSQRTSS (SWFeatureLevel.h:89-116 fastRcp/fastRSqrt/fastDiv) into lookup
 * tables of 1 << *rcp_bits and 2 << *rsqrt_bits entries (buffers sized for max_bits <= 16). */
OSB200_API int osb200_calibrate_sse(uint32_t *rcp_table, int *rcp_bits, uint32_t *rsqrt_table,
                                    int *rsqrt_bits, int max_bits);
/* cudaStream_t the context launches on (for callers that time with events) */
OSB200_API void *osb200_stream(osb200_ctx *ctx);
/* kernels launched by this context since creation (bench.py's gpu_launches) */
OSB200_API uint64_t osb200_launch_count(osb200_ctx *ctx);

/* Per-kernel device time, measured with CUDA events recorded on the context stream around every launch
 * while timing is enabled (bench.py's roofline numbers).  osb200_timing_read synchronises the stream,
 * fills up to `cap` records (one per kernel class, fixed order), resets the accumulators and returns
 * the number of records written. */
typedef struct osb200_kernel_time {
	const char *name; /* static string, e.g. "k_build_lines" */
	uint64_t launches;
	double total_ms;
} osb200_kernel_time;
OSB200_API int osb200_timing_enable(osb200_ctx *ctx, int on);
OSB200_API int osb200_timing_read(osb200_ctx *ctx, osb200_kernel_time *out, int cap);

/* ---- map residency (replaces SWRenderer::SetGameMap + SWMapRenderer ctor's RLE build,
 *      SWRenderer.cpp:172-191, SWMapRenderer.cpp:107-203) ----
 * solid: uint64[512][512] indexed [x][y], bit z set = solid (GameMap.h:160)
 * color: uint32[512][512][64] indexed [x][y][z], 0xHHBBGGRR (GameMap.h:161)            */
OSB200_API int osb200_map_upload(osb200_ctx *ctx, const uint64_t *solid, const uint32_t *color);
/* SWRenderer::SetGameMap(null) (SWRenderer.cpp:172-181): no map resident any more; buffers are kept */
OSB200_API int osb200_map_clear(osb200_ctx *ctx);
/* replaces SWRenderer::GameMapChanged -> SWFlatMapRenderer::SetNeedsUpdate (SWRenderer.cpp:1148-1154,
 * SWFlatMapRenderer.cpp:107-111): O(1), host-only. */
OSB200_API int osb200_map_mark_dirty(osb200_ctx *ctx, int x, int y, int z);
/* replaces SWFlatMapRenderer::Update -> SWMapRenderer::UpdateRle x5 (SWFlatMapRenderer.cpp:52-92):
 * re-packs and uploads only the dirty columns (and their 4 neighbours' derived data) from the
 * caller's live arrays.  Returns the number of columns uploaded (>= 0) or < 0. */
OSB200_API int osb200_map_flush(osb200_ctx *ctx, const uint64_t *solid, const uint32_t *color);
/* read back the device-resident solid map ([x][y] order) -- test hook */
OSB200_API int osb200_map_read_solid(osb200_ctx *ctx, uint64_t *solid);

/* ---- GameMap::CastRay2 (GameMap.cpp:361-507), batched ----
 * v0, dir: n x 3 floats; out: n records.  Host-pointer form copies in/out; the _device form takes
 * device pointers and only enqueues on the context stream. */
OSB200_API int osb200_castray2(osb200_ctx *ctx, const float *v0, const float *dir, int maxSteps,
                               osb200_rayhit *out, size_t n);
OSB200_API int osb200_castray2_device(osb200_ctx *ctx, const float *d_v0, const float *d_dir,
                                      int maxSteps, osb200_rayhit *d_out, size_t n);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * osb200_castray2 (host pointers in and out), but the call allocates nothing: the device staging of the batch and the
 * kernel's work cursor live in a static device array that every GPU loading the library holds until exit (4 MiB + 256 B).
 * n > OSB200_CASTRAY2_STATIC_MAX_RAYS is rejected (-1, text via osb200_last_error).  The array is shared by all calls on a
 * GPU: calls that overlap there (other contexts, streams or host threads) race -- serialise them. */
#define OSB200_CASTRAY2_STATIC_MAX_RAYS 65536
OSB200_API int osb200_castray2_static(osb200_ctx *ctx, const float *v0, const float *dir, int maxSteps,
                                      osb200_rayhit *out, size_t n);

/* ---- terrain pass + fog: SWRenderer::StartScene + EndScene for the map (SWRenderer.cpp:549-563,709-745;
 *      SWMapRenderer::Render SWMapRenderer.cpp:1176-1194; SWRenderer::ApplyFog SWRenderer.cpp:341-516) ----
 * Renders nViews independent scenes of w x h pixels (both multiples of 8, SWRenderer.cpp:105-107).
 * bgrx: nViews*w*h uint32, memory order B,G,R,X, top-down rows (== Bitmap pixels of SWPort::GetFramebuffer);
 * depth: nViews*w*h float view-space Z (== SWRenderer::depthBuffer).  The depth buffer is read-modify:
 * like the reference it is not cleared, and a view whose eye is inside a solid voxel keeps its previous
 * depth and gets the clear colour 0x7f7f7f before fog (SWMapRenderer.cpp:1181-1184).
 * Without a map (SetGameMap(null)) every view takes that same path: clear colour fogged over the stale depth,
 * exactly what SWRenderer::EndScene does when mapRenderer is null (SWRenderer.cpp:713-744).
 * osb200_render keeps context-owned device buffers (persistent depth) and copies to the nullable host
 * pointers, then synchronises; osb200_render_device writes caller-owned device buffers and only enqueues. */
OSB200_API int osb200_render(osb200_ctx *ctx, const osb200_scene *scenes, int nViews, int w, int h,
                             const float fogColor[3], float fogDist, uint32_t *host_bgrx, float *host_depth);
OSB200_API int osb200_render_device(osb200_ctx *ctx, const osb200_scene *scenes, int nViews, int w, int h,
                                    const float fogColor[3], float fogDist, uint32_t *d_bgrx, float *d_depth);

/* SWRenderer::depthBuffer of the frames last rendered by osb200_render (context-owned, persistent): the first n floats */
OSB200_API int osb200_read_depth(osb200_ctx *ctx, float *out, size_t n);

/* ---- sun-shadow bitmap + radiosity: the CPU halves of GLMapShadowRenderer / GLRadiosityRenderer ----
 * State is created lazily by the first call below and starts like freshly constructed reference objects:
 * every shadow pixel and every 16^3 chunk dirty, volumes filled with 0x20080200 (GLRadiosityRenderer.cpp:79-88).
 * Edits reach it through osb200_map_mark_dirty (= GLMapShadowRenderer::GameMapChanged, GLMapShadowRenderer.cpp:220-229)
 * followed by osb200_map_flush; each call here first runs GLMapShadowRenderer::Update (:78-108) for the pending marks,
 * which regenerates the marked 32-pixel words and invalidates the radiosity chunks around every changed pixel
 * (GLRadiosityRenderer::GameMapChanged/Invalidate, GLRadiosityRenderer.cpp:277-336).  Calling with unflushed
 * edits is an error (-1). */

/* replaces GLMapShadowRenderer::Update + read of its `bitmap` (GLMapShadowRenderer.cpp:78-218):
 * out = uint32[512*512] indexed [y*512 + x]: depth<<24 | B6<<16 | G6<<8 | side<<7 | R6.  out may be NULL. */
OSB200_API int osb200_shadow_bitmap(osb200_ctx *ctx, uint32_t *out);
/* replaces GLRadiosityRenderer::Update/UpdateDirtyChunks/UpdateChunk (GLRadiosityRenderer.cpp:348-450,517-543):
 * evaluates up to maxChunks dirty chunks (< 0: all of them) and stores Evaluate + EncodeValue x4 for every voxel
 * of their dirty boxes.  eye (nullable, 3 floats = SceneDefinition::viewOrigin) gives the reference's priority:
 * dirty chunks within 6 chunks of the eye first, any dirty chunk only when there is none (:398-430); among the
 * candidates the reference picks 8 at random, this picks the first maxChunks in chunk-index order.
 * Returns the number of chunks updated (>= 0) or < 0. */
OSB200_API int osb200_radiosity_update(osb200_ctx *ctx, int maxChunks, const float *eye);
/* chunk (cx 0..31, cy 0..31, cz 0..3): out = {dirty, minX, minY, minZ, maxX, maxY, maxZ} (GLRadiosityRenderer.h:52-59;
 * the box is meaningful only when dirty) */
OSB200_API int osb200_radiosity_chunk_info(osb200_ctx *ctx, int cx, int cy, int cz, int out[7]);
/* Chunk::dataFlat/dataX/dataY/dataZ (GLRadiosityRenderer.h:46-51): 4096 RGB10A2 words each, [z][y][x]; any may be NULL */
OSB200_API int osb200_radiosity_read_chunk(osb200_ctx *ctx, int cx, int cy, int cz, uint32_t *flat, uint32_t *x,
                                           uint32_t *y, uint32_t *z);
/* GLRadiosityRenderer::Evaluate (GLRadiosityRenderer.cpp:162-275) un-encoded over the inclusive voxel box:
 * 12 floats per voxel (base.rgb, x.rgb, y.rgb, z.rgb), box-ordered [z][y][x].  Coordinates may run past the
 * map edge in x / y (the bitmap wraps).  Independent of the chunk state -- parity hook. */
OSB200_API int osb200_radiosity_evaluate_box(osb200_ctx *ctx, int x0, int y0, int z0, int x1, int y1, int z1,
                                             float *out);
/* device pointers of the four resident 512x512x64 RGB10A2 volumes ([z][y][x], the layout of the reference's
 * 3-D textures, GLRadiosityRenderer.cpp:120-129) and of the 512x512 shadow bitmap; any may be NULL */
/* Measurement hook: number of (voxel, tap) pairs of a full-map Evaluate that pass the normal cull
 * (GLRadiosityRenderer.cpp:199-214) and execute the float work; total = 512*512*64*169.  Used for the FP32 figure of C5. */
OSB200_API int osb200_radiosity_count_taps(osb200_ctx *ctx, unsigned long long *executed, unsigned long long *total);
OSB200_API int osb200_radiosity_volumes(osb200_ctx *ctx, const uint32_t **flat, const uint32_t **x,
                                        const uint32_t **y, const uint32_t **z, const uint32_t **shadow);

/* ---- the full EndScene of the software renderer: terrain, voxel models, dynamic lights, fog (SWRenderer.cpp:709-745) ----
 * A model is created once from a VoxelModel's arrays (= SWRenderer::CreateModel -> SWModel::SWModel, SWModel.cpp:25-100):
 * solid[x + y*w] has bit z set for a solid voxel, colors[(x + y*w)*d + z] = material << 24 | 0xBBGGRR as VoxelModel
 * stores them (VoxelModel.h:68-106), origin = VoxelModel::GetOrigin().  d <= 64. */
typedef struct osb200_model osb200_model;
OSB200_API int osb200_model_create(osb200_ctx *ctx, int w, int h, int d, const float origin[3], const uint64_t *solid,
                                   const uint32_t *colors, osb200_model **out);
OSB200_API void osb200_model_destroy(osb200_model *model);
/* SWModel::renderData / renderDataAddr (test hook): returns renderData.size(); copies when the capacities suffice */
OSB200_API int osb200_model_render_data(const osb200_model *model, uint32_t *data, int dataCap, uint32_t *addr, int addrCap);

/* one IRenderer::RenderModel call: client::ModelRenderParam's matrix (column-major, Matrix4::m) and customColor
 * (IRenderer.h:39-63; depthHack / ghost / opacity are not read by the software renderer) */
typedef struct osb200_model_instance {
	const osb200_model *model;
	float matrix[16];
	float customColor[3];
} osb200_model_instance;
/* one IRenderer::AddLight call; only point lights are drawn by the software renderer (SWRenderer.cpp:581-585) */
typedef struct osb200_light {
	float origin[3];
	float radius;
	float color[3];
} osb200_light;
/* StartScene + RenderModel x nInst + AddLight x nLights + EndScene for one view: terrain (as osb200_render), then every
 * model splatted voxel by voxel with the reference's strict depth test and submission order (SWModelRenderer.cpp:46-277),
 * then the dynamic lights in submission order (SWRenderer::AddLight culls :578-649, ApplyDynamicLight :229-339), then the
 * fog.  Frame and depth live in the context (persistent depth, see osb200_render) and are copied to the nullable host
 * pointers; the call synchronises. */
OSB200_API int osb200_render_scene(osb200_ctx *ctx, const osb200_scene *scene, int w, int h, const float fogColor[3],
                                   float fogDist, const osb200_model_instance *inst, int nInst, const osb200_light *lights,
                                   int nLights, uint32_t *host_bgrx, float *host_depth);

/* ---- VXL stream <-> resident map on the device.
 *      osb200_map_load_vxl replaces GameMap::Load (Client/GameMap.cpp:519-596) + osb200_map_upload: the raw (inflated)
 *      VXL bytes go to the device (2 MB for Title.vxl instead of 66 MiB of arrays), column starts are found by pointer
 *      doubling and every column is decoded by its own thread into the resident map.  host_solid / host_color
 *      (nullable) receive the arrays in GameMap's [x][y] / [x][y][z] order -- what the owner of the map keeps for
 *      GameMap::Set and osb200_map_flush.  Truncated / malformed input: -1 (the reference throws), no map loaded.
 *      osb200_map_save_vxl replaces GameMap::Save (:93-160, IsSurface :65-83): returns the stream length and writes
 *      min(length, cap) bytes to out (out may be NULL with cap 0 to ask for the length). ---- */
OSB200_API int osb200_map_load_vxl(osb200_ctx *ctx, const uint8_t *data, size_t len, uint64_t *host_solid,
                                   uint32_t *host_color);
OSB200_API long long osb200_map_save_vxl(osb200_ctx *ctx, uint8_t *out, size_t cap);

/* ---- Floating-block detection: GameMapWrapper (Client/GameMapWrapper.cpp) on the resident map.  State: one "linked"
 *      bit per voxel (connected to the ground through solid 6-neighbours, no wrap; the z = 63 layer is the root).
 *      All three calls need a flushed map (osb200_map_flush).
 *      osb200_links_rebuild        GameMapWrapper::Rebuild (:50-102); returns the number of flood launches (>= 1).
 *      osb200_links_blocks_added   after voxels were added (GameMapWrapper::AddBlock, :104-193) and flushed.
 *      osb200_links_remove_blocks  GameMapWrapper::RemoveBlocks (:199-361) for cells {x, y, z} x n (int16) that the
 *                                  caller has already cleared on the host map and flushed (the reference's
 *                                  m.Set(.., false) at :214): returns how many solid voxels lost their connection
 *                                  and writes min(count, cap) of them, sorted by (x, y, z) (the reference returns
 *                                  the same set in its walk order).  Like the reference it does not remove them.
 *      osb200_links_read           linked bitmap in [x][y] order, bit z. ---- */
OSB200_API int osb200_links_rebuild(osb200_ctx *ctx);
OSB200_API int osb200_links_blocks_added(osb200_ctx *ctx);
OSB200_API int osb200_links_remove_blocks(osb200_ctx *ctx, const int16_t *cells, int n, int16_t *floating, int cap);
OSB200_API int osb200_links_read(osb200_ctx *ctx, uint64_t *out);
/* the result of the last osb200_links_remove_blocks again (for a caller whose buffer was too small): count, min(count, cap) written */
OSB200_API int osb200_links_last_floating(osb200_ctx *ctx, int16_t *floating, int cap);
OSB200_API int osb200_links_last_rounds(osb200_ctx *ctx); /* flood launches of the last links call (diagnostics) */

/* ---- Falling objects: World::ClusterizeBlocks (Client/World.cpp:154-226), the step World::ApplyBlockActions
 *      (:227-265) runs on RemoveBlocks' result: the 6-connected components of a set of cells {x, y, z} x n (int16,
 *      inside the map).  cluster[i] = number of the component of cell i; components are numbered 0 .. count-1 by
 *      their smallest cell in CellPos order (the reference returns them in the iteration order of its
 *      unordered_map: the same partition).  A cell given twice is one cell.  Needs no map.  Returns count. ---- */
OSB200_API int osb200_clusterize_blocks(osb200_ctx *ctx, const int16_t *cells, int n, int32_t *cluster);

/* ---- Minimap: SWFlatMapRenderer::Update / GeneratePixel (Draw/SWFlatMapRenderer.cpp:52-105): colour of the topmost solid
 *      voxel of every column, 0xffRRGGBB-swizzled like the reference's SWImage, out[y * 512 + x] (nullable); *device
 *      (nullable) receives the device copy.  Needs a flushed map. ---- */
OSB200_API int osb200_minimap(osb200_ctx *ctx, uint32_t *out, const uint32_t **device);

/* ---- GameMap::CastRay (GameMap.cpp:208-359), batched: the integer cell-wall walk used by the ambient-occlusion
 *      evaluation and by the game's hit tests.  v0, dir: n x 3 floats; one length for the batch (the walk takes at most
 *      (long)length cell walls); out: n x {hit, hitBlock.x, hitBlock.y, hitBlock.z} int32 (zeros on a miss). ---- */
OSB200_API int osb200_castray(osb200_ctx *ctx, const float *v0, const float *dir, float length, int32_t *out, size_t n);
OSB200_API int osb200_castray_device(osb200_ctx *ctx, const float *d_v0, const float *d_dir, float length,
                                     int32_t *d_out, size_t n);

/* ---- ray-traced ambient occlusion: the CPU half of GLAmbientShadowRenderer (Draw/GLAmbientShadowRenderer.cpp) ----
 * State is created lazily and starts like a freshly constructed reference object: every 16^3 chunk dirty over its whole
 * box, the volume filled with (1, 1).  Edits reach it through osb200_map_mark_dirty (= GameMapChanged -> Invalidate of
 * the +-16 box, :171-237) followed by osb200_map_flush; updating with unflushed edits is an error (-1).
 * The 16 ray directions are an input: the reference draws them per instance from its global RNG (:53-60), so a caller
 * that wants the reference's values passes the reference's rays; the defaults follow the same recipe from a fixed seed. */
OSB200_API int osb200_ambient_set_rays(osb200_ctx *ctx, const float rays[48]);
OSB200_API int osb200_ambient_get_rays(osb200_ctx *ctx, float rays[48]);
/* replaces GLAmbientShadowRenderer::Update/UpdateDirtyChunks/UpdateChunk (:239-507): re-evaluates up to maxChunks dirty
 * chunks (< 0: all): Evaluate (16 rays x GameMap::CastRay, :130-169) over the dirty box grown by 2, the under-shadowing
 * compensation, two rounds of in-place 3-tap blurs along x, y, z, copy-out.  eye (nullable) gives the reference's
 * priority: dirty chunks within 6 chunks of the eye first (:276-292); among the candidates the reference draws 8 at
 * random, this takes the first maxChunks in chunk-index order.  Returns the number of chunks updated or < 0. */
OSB200_API int osb200_ambient_update(osb200_ctx *ctx, int maxChunks, const float *eye);
/* out = {dirty, minX, minY, minZ, maxX, maxY, maxZ} (GLAmbientShadowRenderer.h:55-62) */
OSB200_API int osb200_ambient_chunk_info(osb200_ctx *ctx, int cx, int cy, int cz, int out[7]);
/* Chunk::data (GLAmbientShadowRenderer.h:54): float[16][16][16][2] = [z][y][x]{AO term, 1 for air / 0 for solid} */
OSB200_API int osb200_ambient_read_chunk(osb200_ctx *ctx, int cx, int cy, int cz, float *out);
/* GLAmbientShadowRenderer::Evaluate over the inclusive voxel box, [z][y][x]; coordinates may leave the map (x / y wrap,
 * z < 0 is air, z >= 64 solid).  Independent of the chunk state -- parity hook. */
OSB200_API int osb200_ambient_evaluate_box(osb200_ctx *ctx, int x0, int y0, int z0, int x1, int y1, int z1, float *out);
/* device pointer of the resident volume in the layout of the reference's RG-float 3-D texture (:93-105,255-262):
 * float2[65][512][512], slice z + 1 holds map depth z, slice 0 stays (1, 1) */
OSB200_API int osb200_ambient_volume(osb200_ctx *ctx, const float **d_volume);


/* ---- several GPUs of one box from one process (the reference has nothing distributed) ----
 * A group owns one context per listed device and a NCCL communicator over them (libnccl.so.2 is bound at run time and
 * only when nDevices > 1).  The map is replicated: uploaded over PCIe once and copied to the peers with ncclBroadcast
 * over NVLink; edits likewise (the dirty columns are uploaded once, their records broadcast, every device scatters them).
 * Views and rays are independent units: they are sharded by contiguous range [i*n/G, (i+1)*n/G) with no collective on
 * the data path, one host thread per device.  Results are bit-identical to a single-device call on the same inputs. */
typedef struct osb200_group osb200_group;
OSB200_API int osb200_group_create(const osb200_opts *opts /* device field ignored */, const int *devices, int nDevices,
                                   osb200_group **out);
OSB200_API void osb200_group_destroy(osb200_group *g);
OSB200_API int osb200_group_size(const osb200_group *g);
OSB200_API osb200_ctx *osb200_group_ctx(osb200_group *g, int i); /* for the per-device calls (radiosity, timing ...) */
OSB200_API int osb200_group_map_upload(osb200_group *g, const uint64_t *solid, const uint32_t *color);
OSB200_API int osb200_group_map_mark_dirty(osb200_group *g, int x, int y, int z);
/* returns the number of columns flushed (on every device) */
OSB200_API int osb200_group_map_flush(osb200_group *g, const uint64_t *solid, const uint32_t *color);
OSB200_API int osb200_group_render(osb200_group *g, const osb200_scene *scenes, int nViews, int w, int h,
                                   const float fogColor[3], float fogDist, uint32_t *host_bgrx, float *host_depth);
/* same views, the finished frames gathered on device `root` over NVLink (grouped ncclSend / ncclRecv: only the root
 * receives); d_bgrx_root / d_depth_root (nullable) are device pointers on the root's device, nViews * w * h words */
OSB200_API int osb200_group_render_gather(osb200_group *g, const osb200_scene *scenes, int nViews, int w, int h,
                                          const float fogColor[3], float fogDist, int root, uint32_t *d_bgrx_root,
                                          float *d_depth_root);
OSB200_API int osb200_group_castray2(osb200_group *g, const float *v0, const float *dir, int maxSteps, osb200_rayhit *out,
                                     size_t n);

#ifdef __cplusplus
}
#endif
#endif /* OSB200_H */
