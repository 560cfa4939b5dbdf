/*
 * osb_oracle -- CPU restatement ("port") of the reference algorithms on the hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this; the product (openspades_b200/, include/) never does.
 *
 * Parity pinning: the reference (yvt/openspades) ships no tests, golden vectors or KATs for this
 * path, so this restatement is pinned against the UNMODIFIED reference code
 * itself, compiled in place into oracle/_ref/liboref_{sse2,ieee}.so (oracle/Makefile) and run on
 * the same inputs: tests/test_oracle_vs_ref.py, plus the committed fixtures under tests/golden/
 * that were generated from those libraries by tests/golden/make_golden.py.
 *
 * Every function cites the reference file:line it follows (paths relative to Sources/).
 */
#ifndef OSB_ORACLE_H
#define OSB_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_scene { /* same layout as osb200_scene / oref_scene */
	float origin[3];
	float axis[3][3];
	float fovX, fovY;
	float zNear, zFar;
} orc_scene;

typedef struct orc_rayhit { /* GameMap::RayCastResult, Client/GameMap.h:150-156 */
	uint8_t hit, startSolid, pad[2];
	float hitPos[3];
	int32_t hitBlock[3];
	int32_t normal[3];
} orc_rayhit;

enum { ORC_NUM_IEEE = 0, ORC_NUM_SSE = 1, ORC_NUM_SSE2 = 2 };

/* map arrays everywhere: solid uint64[512][512] ([x][y]), color uint32[512][512][64] ([x][y][z]) */

/* Client/GameMap.cpp:519-596 GameMap::Load.  Returns 0 or <0 on truncated input. */
int orc_vxl_load(const uint8_t *data, size_t len, uint64_t *solid, uint32_t *color);
/* Client/GameMap.cpp:93-160 GameMap::Save (IsSurface :65-83).  Returns the stream length, writes <= cap bytes. */
long long orc_vxl_save(const uint64_t *solid, const uint32_t *color, uint8_t *out, size_t cap);

/* Client/GameMap.cpp:361-507 GameMap::CastRay2 */
void orc_castray2(const uint64_t *solid, const float *v0, const float *dir, int maxSteps, orc_rayhit *out,
                  size_t n, int threads);

/* Draw/SWMapRenderer.cpp:142-203 BuildRle: writes the byte image of the column's RLE record,
 * returns its length (multiple of 4). out must hold >= 400 bytes. */
int orc_build_rle(const uint64_t *solid, int x, int y, int8_t *out);

/* Draw/SWRenderer.cpp:709-745 EndScene (clear, terrain, fog) for one view.
 * numerics: ORC_NUM_*; threads = r_swNumThreads; under = r_swUndersampling clamp(1,2,4).
 * bgrx/depth: w*h each; depth is read-modify (the reference never clears it). */
int orc_render(const uint64_t *solid, const uint32_t *color, const orc_scene *scene, int w, int h,
               const float fogColor[3], float fogDist, int numerics, int threads, int under,
               uint32_t *bgrx, float *depth);
/* strip buffer of the last orc_render on this thread (debug): returns numLines */
int orc_last_lines(int *lineRes, const uint32_t **combined, const float **depth);

/* Draw/GLMapShadowRenderer.cpp:177-218 GeneratePixel over the whole map: out uint32[y*512+x] */
void orc_shadow_bitmap(const uint64_t *solid, const uint32_t *color, uint32_t *out);
uint32_t orc_shadow_pixel(const uint64_t *solid, const uint32_t *color, int x, int y);
/* Draw/GLMapShadowRenderer.cpp:220-229 GameMapChanged/MarkUpdate on a 512*16-word dirty bitset */
void orc_shadow_mark(uint32_t *words, int x, int y, int z);
/* Draw/GLMapShadowRenderer.cpp:78-108 Update: regenerate marked words, invalidate radiosity chunk boxes
 * (radState: 4096 x {dirty,minX,minY,minZ,maxX,maxY,maxZ}, index (cx + cy*32)*4 + cz; may be NULL).
 * Returns the number of pixels that changed. */
int orc_shadow_update(const uint64_t *solid, const uint32_t *color, uint32_t *bitmap, uint32_t *words,
                      int32_t *radState);
/* Draw/GLRadiosityRenderer.cpp:277-336 GameMapChanged + Invalidate */
void orc_radiosity_invalidate(int32_t *radState, int x, int y, int z);
/* Draw/GLRadiosityRenderer.cpp:162-275 Evaluate: 12 floats (base,x,y,z as rgb each) */
void orc_radiosity_evaluate(const uint32_t *bitmap, int x, int y, int z, float out[12]);
/* Draw/GLRadiosityRenderer.cpp:456-515 EncodeValue (mode = r_radiosity) */
uint32_t orc_radiosity_encode(const float v[3], int mode);
/* Evaluate + EncodeValue over a box [x0,x1]x[y0,y1]x[z0,z1] (UpdateChunk :517-543 inner loop);
 * f32 (12 floats/voxel) and enc (4 words/voxel: flat,x,y,z) are box-ordered [z][y][x]; either may be NULL */
void orc_radiosity_box(const uint32_t *bitmap, int x0, int y0, int z0, int x1, int y1, int z1, int mode,
                       float *f32, uint32_t *enc, int threads);

/* Client/GameMap.cpp:208-359 GameMap::CastRay.  out[i] = {hit, hitBlock.x, .y, .z} (zeros on a miss) */
int orc_castray_one(const uint64_t *solid, const float v0[3], const float dir[3], float length, int32_t hit[3]);
void orc_castray(const uint64_t *solid, const float *v0, const float *dir, float length, int32_t *out, size_t n,
                 int threads);
/* Draw/GLAmbientShadowRenderer.cpp:130-169 Evaluate (rays = the instance's 16 directions, 48 floats) */
float orc_ao_evaluate(const uint64_t *solid, const float rays[48], int x, int y, int z);
void orc_ao_evaluate_box(const uint64_t *solid, const float rays[48], int x0, int y0, int z0, int x1, int y1, int z1,
                         float *out, int threads);
/* chunk state 4096 x {dirty,minX,minY,minZ,maxX,maxY,maxZ}, index (cx + cy*32)*4 + cz */
void orc_ao_state_init(int32_t *state);
/* Draw/GLAmbientShadowRenderer.cpp:171-237 GameMapChanged + Invalidate */
void orc_ao_invalidate(int32_t *state, int x, int y, int z);
/* Draw/GLAmbientShadowRenderer.cpp:334-507 UpdateChunk; data = the chunk's float[16][16][16][2] */
void orc_ao_update_chunk(const uint64_t *solid, const float rays[48], int32_t *state, int cx, int cy, int cz,
                         float *data);

/* ---- voxel models + dynamic lights inside EndScene (orc_models.c) ---- */
typedef struct orc_model orc_model;
typedef struct orc_model_instance { /* one SWRenderer::RenderModel call: same layout as osb200_model_instance */
	const orc_model *model;
	float matrix[16]; /* Matrix4::m, column-major */
	float customColor[3];
} orc_model_instance;
typedef struct orc_light { /* client::DynamicLightParam of a point light */
	float origin[3];
	float radius;
	float color[3];
} orc_light;
/* Draw/SWModel.cpp:25-100; solid[x + y*w] bit z, colors[(x + y*w)*d + z] */
orc_model *orc_model_create(int w, int h, int d, const float origin[3], const uint64_t *solid, const uint32_t *colors);
void orc_model_free(orc_model *m);
/* SWModel::renderData / renderDataAddr; returns the length of renderData */
int orc_model_render_data(const orc_model *m, uint32_t *data, int dataCap, uint32_t *addr, int addrCap);
/* Draw/SWRenderer.cpp:709-745 EndScene: clear, terrain (skipped when solid == NULL), models, lights, fog */
int orc_render_scene(const uint64_t *solid, const uint32_t *color, const orc_scene *scene, int w, int h,
                     const float fogColor[3], float fogDist, int numerics, int threads, int under,
                     const orc_model_instance *inst, int nInst, const orc_light *lights, int nLights, uint32_t *bgrx,
                     float *depth);

/* ---- floating-block detection + minimap (orc_links.c) ---- */
typedef struct orc_links orc_links;
orc_links *orc_links_create(void);
void orc_links_free(orc_links *w);
/* Client/GameMapWrapper.cpp:50-102 Rebuild */
void orc_links_rebuild(orc_links *w, const uint64_t *solid);
/* :104-193 AddBlock (sets the voxel in the caller's arrays like GameMap::Set) */
void orc_links_add_block(orc_links *w, uint64_t *solid, uint32_t *color, int x, int y, int z, uint32_t col);
/* :199-361 RemoveBlocks: clears the cells, returns the floating blocks (count; min(count, cap) written, reference order) */
int orc_links_remove_blocks(orc_links *w, uint64_t *solid, const int16_t *cells, int n, int16_t *out, int cap);
void orc_links_linked(const orc_links *w, uint64_t *out);
/* Client/World.cpp:154-226 ClusterizeBlocks: cells n x {x, y, z} (0 <= x, y < 4096, 0 <= z < 256); cluster[i] = piece of
 * cell i, pieces numbered in order of their first cell in the input; returns the number of pieces */
int orc_clusterize_blocks(const int16_t *cells, int n, int32_t *cluster);
/* Draw/SWFlatMapRenderer.cpp:94-105 GeneratePixel over the map: out[y * 512 + x] */
void orc_minimap(const uint64_t *solid, const uint32_t *color, uint32_t *out);

#ifdef __cplusplus
}
#endif
#endif
