/*
 * oref -- C entry points into the UNMODIFIED reference code (yvt/openspades), compiled in
 * place from /root/reference by oracle/Makefile into oracle/_ref/liboref_{sse2,ieee}.so.
 *
 * TEST INFRASTRUCTURE ONLY: used by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs as the checker and the timed CPU baseline.  Nothing
 * in the product (openspades_b200/, include/) links, loads or calls it.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OREF_API __attribute__((visibility("default")))

/* Mirrors client::SceneDefinition's fields used by the terrain path (SceneDefinition.h:26-33). */
typedef struct oref_scene {
	float origin[3];
	float axis[3][3]; /* viewAxis[0]=right, [1]=up, [2]=front */
	float fovX, fovY;
	float zNear, zFar;
} oref_scene;

/* Same field order/offsets as GameMap::RayCastResult (GameMap.h:150-156), 40 bytes. */
typedef struct oref_rayhit {
	uint8_t hit, startSolid, pad[2];
	float hitPos[3];
	int32_t hitBlock[3];
	int32_t normal[3];
} oref_rayhit;

OREF_API const char *oref_variant(void); /* "sse2" or "ieee" (compile-time) */

/* ---- GameMap ---- */
OREF_API void *oref_map_load_vxl(const char *path);  /* GameMap::Load, GameMap.cpp:519 */
OREF_API void *oref_map_load_vxl_mem(const void *data, size_t len);
OREF_API long long oref_map_save_vxl(void *map, void *out, size_t cap); /* GameMap::Save, GameMap.cpp:93-160 */
OREF_API void *oref_map_from_arrays(const uint64_t *solid /*[x][y]*/, const uint32_t *color /*[x][y][z]*/);
OREF_API void oref_map_free(void *map);
OREF_API void oref_map_get_arrays(void *map, uint64_t *solid, uint32_t *color);
/* GameMapWrapper (Client/GameMapWrapper.cpp): create = ctor + Rebuild */
OREF_API void *oref_wrapper_create(void *map);
OREF_API void oref_wrapper_free(void *w);
OREF_API void oref_wrapper_add_block(void *w, int x, int y, int z, uint32_t color);
OREF_API int oref_wrapper_remove_blocks(void *w, const int16_t *cells, int n, int16_t *out, int cap);
OREF_API void oref_wrapper_linked(void *w, uint64_t *out /*[x][y], bit z = link != Invalid*/);
/* World::ClusterizeBlocks (Client/World.cpp:154-226, oref_world.cpp): cluster[i] = index of the returned piece holding cell i */
OREF_API int oref_clusterize_blocks(const int16_t *cells, int n, int32_t *cluster);
OREF_API void oref_map_set(void *map, int x, int y, int z, int solid, uint32_t color); /* GameMap::Set, notifies listeners */
OREF_API void oref_castray2(void *map, const float *v0, const float *dir, int maxSteps,
                            oref_rayhit *out, size_t n, int threads); /* GameMap::CastRay2 */

/* ---- SWRenderer ---- */
/* level: -1 = DetectFeatureLevel(), else SWFeatureLevel enum value of this build */
OREF_API void *oref_renderer_create(int w, int h, int level, int threads);
OREF_API void oref_renderer_free(void *r);
OREF_API int oref_renderer_level(void *r);
OREF_API void oref_renderer_set_threads(int threads); /* cvar r_swNumThreads */
OREF_API void oref_renderer_set_undersampling(int under); /* cvar r_swUndersampling, 0 = off */
OREF_API void oref_renderer_set_map(void *r, void *map);
OREF_API int oref_renderer_minimap(void *r, uint32_t *out); /* SWFlatMapRenderer image, [y][x] */
OREF_API void oref_renderer_set_fog(void *r, float cr, float cg, float cb, float dist);
/* StartScene + EndScene + FrameDone; copies framebuffer (BGRX) and depth when non-NULL.
   Returns 0, or -1 when the reference threw. */
OREF_API int oref_renderer_render(void *r, const oref_scene *s, uint32_t *bgrx, float *depth);
OREF_API const char *oref_last_error(void);
/* raw RLE bytes of column (x,y) as SWMapRenderer stores them; returns length */
OREF_API int oref_renderer_rle(void *r, int x, int y, int8_t *out, int cap);
/* strip buffer of the last frame (for debugging parity): numLines, lineResolution */
OREF_API int oref_renderer_lines(void *r, int *lineRes);
OREF_API int oref_renderer_line(void *r, int i, uint32_t *combined, float *depth, float *misc /*horizX,horizY,pitchTanMin,pitchScale*/, int32_t *misci);

/* ---- voxel models + dynamic lights inside EndScene (SWModel, SWModelRenderer, SWRenderer::AddLight / ApplyDynamicLight) ---- */
typedef struct oref_model_param { /* client::ModelRenderParam fields the SW path reads (IRenderer.h:39-63) */
	float matrix[16];
	float customColor[3];
} oref_model_param;
typedef struct oref_light { /* client::DynamicLightParam, point lights only (SWRenderer.cpp:581-585) */
	float origin[3];
	float radius;
	float color[3];
} oref_light;
OREF_API void *oref_vmodel_load_kv6(const void *data, size_t len);   /* VoxelModel::LoadKV6 */
OREF_API void *oref_vmodel_from_arrays(int w, int h, int d, const float origin[3], const uint64_t *solid /*[x + y*w]*/,
                                       const uint32_t *colors /*[(x + y*w)*d + z]*/);
OREF_API void oref_vmodel_free(void *vm);
OREF_API void oref_vmodel_info(void *vm, int dims[3], float origin[3]);
OREF_API void oref_vmodel_arrays(void *vm, uint64_t *solid, uint32_t *colors);
OREF_API void *oref_model_create(void *renderer, void *vm);           /* SWRenderer::CreateModel -> SWModel ctor */
OREF_API void oref_model_free(void *model);
/* SWModel::renderData / renderDataAddr (SWModel.cpp:25-100): returns renderData.size(); copies when the caps suffice */
OREF_API int oref_model_render_data(void *model, uint32_t *data, int dataCap, uint32_t *addr, int addrCap);
/* StartScene + RenderModel x nModels + AddLight x nLights + EndScene + FrameDone */
OREF_API int oref_renderer_render_scene(void *r, const oref_scene *s, void *const *models, const oref_model_param *params,
                                        int nModels, const oref_light *lights, int nLights, uint32_t *bgrx, float *depth);

/* ---- CPU halves of GLMapShadowRenderer / GLRadiosityRenderer (oref_gl.cpp) ---- */
OREF_API void *oref_rad_create(void *map, int mode /* r_radiosity: 1 sqrt-compressed, 2 linear */);
OREF_API void oref_rad_free(void *h);
OREF_API void oref_rad_map_changed(void *h, int x, int y, int z);  /* GLMapShadowRenderer::GameMapChanged */
OREF_API void oref_rad_shadow_update(void *h);                     /* GLMapShadowRenderer::Update */
OREF_API void oref_rad_shadow_bitmap(void *h, uint32_t *out /*[y*512+x]*/);
OREF_API uint32_t oref_rad_generate_pixel(void *h, int x, int y);   /* GLMapShadowRenderer::GeneratePixel */
OREF_API void oref_rad_evaluate(void *h, int x, int y, int z, float out[12]); /* GLRadiosityRenderer::Evaluate */
OREF_API void oref_rad_evaluate_box(void *h, int x0, int y0, int z0, int x1, int y1, int z1, float *out, int threads);
OREF_API uint32_t oref_rad_encode(void *h, const float v[3]);      /* GLRadiosityRenderer::EncodeValue */
OREF_API void oref_rad_chunk_info(void *h, int cx, int cy, int cz, int out[7]);
OREF_API int oref_rad_num_dirty(void *h);
OREF_API void oref_rad_update_chunk(void *h, int cx, int cy, int cz); /* GLRadiosityRenderer::UpdateChunk */
OREF_API int oref_rad_update_all(void *h, int threads);
OREF_API void oref_rad_read_chunk(void *h, int cx, int cy, int cz, uint32_t *flat, uint32_t *x, uint32_t *y, uint32_t *z);

/* ---- GameMap::CastRay (GameMap.cpp:208-359) and the CPU half of GLAmbientShadowRenderer (oref_gl.cpp) ---- */
/* out[i] = {hit, hitBlock.x, hitBlock.y, hitBlock.z}; hitBlock is left 0 on a miss */
OREF_API void oref_castray(void *map, const float *v0, const float *dir, float length, int32_t *out, size_t n, int threads);
OREF_API void *oref_ao_create(void *map);
OREF_API void oref_ao_free(void *h);
OREF_API void oref_ao_get_rays(void *h, float out[48]);  /* the 16 random directions of this instance */
OREF_API void oref_ao_set_rays(void *h, const float in[48]);
OREF_API float oref_ao_evaluate(void *h, int x, int y, int z); /* GLAmbientShadowRenderer::Evaluate */
OREF_API void oref_ao_evaluate_box(void *h, int x0, int y0, int z0, int x1, int y1, int z1, float *out, int threads);
OREF_API void oref_ao_map_changed(void *h, int x, int y, int z); /* GLAmbientShadowRenderer::GameMapChanged */
OREF_API void oref_ao_chunk_info(void *h, int cx, int cy, int cz, int out[7]);
OREF_API int oref_ao_num_dirty(void *h);
OREF_API void oref_ao_update_chunk(void *h, int cx, int cy, int cz); /* GLAmbientShadowRenderer::UpdateChunk */
OREF_API int oref_ao_update_all(void *h, int threads);
OREF_API void oref_ao_read_chunk(void *h, int cx, int cy, int cz, float *out /*[16][16][16][2]*/);

#ifdef __cplusplus
}
#endif
