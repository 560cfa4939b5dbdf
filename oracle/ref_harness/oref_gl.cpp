/*
 * TEST INFRASTRUCTURE ONLY -- see oref.h.  Drives the CPU halves of the reference's OpenGL renderer
 * (Draw/GLMapShadowRenderer.cpp, Draw/GLRadiosityRenderer.cpp, compiled unmodified and in place)
 * without any graphics API:
 *   - the abstract draw::IGLDevice is satisfied by a generated do-nothing NullGLDevice
 *     (gen_null_gl.py): those two classes use GL only to upload textures;
 *   - they take a GLRenderer& but touch only its inline accessors (GLRenderer.h:226-246) and the
 *     mapShadowRenderer member, so a GLRenderer is never constructed (GLRenderer.cpp and its ~100
 *     GL passes are not linked): the harness hands them zeroed storage in which only `device`,
 *     `settings`, `mapShadowRenderer` and `radiosityRenderer` are brought to life;
 *   - GLProfiler::Context (a GL timer-query scope, GLProfiler.cpp:590-615) is stubbed inactive.
 * This TU is compiled with -fno-access-control to reach private members (bitmap, chunks, Evaluate...).
 */
#include "oref.h"

#include <cstring>
#include <new>
#include <string>

#include <Client/GameMap.h>
#include <Core/Settings.h>
#include <Draw/GLAmbientShadowRenderer.h>
#include <Draw/GLMapShadowRenderer.h>
#include <Draw/GLProfiler.h>
#include <Draw/GLRadiosityRenderer.h>
#include <Draw/GLRenderer.h>

#include "NullGLDevice.h"

using namespace spades;
using namespace spades::draw;

namespace spades {
	namespace draw {
		GLProfiler::Context::Context(GLProfiler &profiler, const char *, ...)
		    : m_profiler{profiler}, m_active{false} {}
		GLProfiler::Context::~Context() {}
	} // namespace draw
} // namespace spades

namespace {
	struct Ao {
		alignas(64) unsigned char storage[sizeof(GLRenderer)];
		GLRenderer *renderer;
		client::GameMap *map;
		GLAmbientShadowRenderer *ao;
	};
	struct Rad {
		alignas(64) unsigned char storage[sizeof(GLRenderer)];
		GLRenderer *renderer;
		client::GameMap *map;
		GLMapShadowRenderer *shadow;
		GLRadiosityRenderer *rad;
	};
} // namespace

extern "C" {

/* mode = r_radiosity (1: sqrt-compressed encode, 2: linear), latched by GLSettings at construction */
void *oref_rad_create(void *map, int mode) {
	Settings::ItemHandle("r_radiosity", nullptr) = mode;
	Rad *h = new Rad();
	std::memset(h->storage, 0, sizeof(h->storage));
	h->renderer = reinterpret_cast<GLRenderer *>(h->storage);
	h->map = static_cast<client::GameMap *>(map);
	new (&h->renderer->device) Handle<IGLDevice>(Handle<oref::NullGLDevice>::New().Cast<IGLDevice>());
	new (&h->renderer->settings) GLSettings();
	h->renderer->map = h->map;
	h->shadow = new GLMapShadowRenderer(*h->renderer, h->map);
	h->renderer->mapShadowRenderer = h->shadow;
	h->rad = new GLRadiosityRenderer(*h->renderer, h->map);
	h->renderer->radiosityRenderer = h->rad;
	h->shadow->Update(); /* initial state: every pixel dirty (GLMapShadowRenderer.cpp:64-65) */
	return h;
}

void oref_rad_free(void *p) {
	Rad *h = static_cast<Rad *>(p);
	delete h->rad;
	delete h->shadow;
	h->renderer->settings.~GLSettings();
	h->renderer->device.~Handle<IGLDevice>();
	delete h;
}

/* GLRenderer::GameMapChanged's shadow-map leg (GLRenderer.cpp:1363-1374) */
void oref_rad_map_changed(void *p, int x, int y, int z) {
	Rad *h = static_cast<Rad *>(p);
	h->shadow->GameMapChanged(x, y, z, h->map);
}

/* GLMapShadowRenderer::Update (:78-175): regenerates dirty 32-pixel words, invalidates radiosity */
void oref_rad_shadow_update(void *p) { static_cast<Rad *>(p)->shadow->Update(); }

void oref_rad_shadow_bitmap(void *p, uint32_t *out) {
	Rad *h = static_cast<Rad *>(p);
	std::memcpy(out, h->shadow->bitmap.data(), h->shadow->bitmap.size() * 4);
}

/* GLMapShadowRenderer::GeneratePixel (:199-218) */
uint32_t oref_rad_generate_pixel(void *p, int x, int y) { return static_cast<Rad *>(p)->shadow->GeneratePixel(x, y); }

/* GLRadiosityRenderer::Evaluate (:162-275): base, x, y, z as 12 floats */
void oref_rad_evaluate(void *p, int x, int y, int z, float out[12]) {
	GLRadiosityRenderer::Result r = static_cast<Rad *>(p)->rad->Evaluate(IntVector3::Make(x, y, z));
	const Vector3 *v[4] = {&r.base, &r.x, &r.y, &r.z};
	for (int i = 0; i < 4; i++) {
		out[i * 3 + 0] = v[i]->x;
		out[i * 3 + 1] = v[i]->y;
		out[i * 3 + 2] = v[i]->z;
	}
}

/* Evaluate over a box, [z][y][x] order, OpenMP over z*y rows (Evaluate only reads) */
void oref_rad_evaluate_box(void *p, int x0, int y0, int z0, int x1, int y1, int z1, float *out, int threads) {
	Rad *h = static_cast<Rad *>(p);
	const int nx = x1 - x0 + 1, ny = y1 - y0 + 1, nz = z1 - z0 + 1;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic, 4)
	for (int row = 0; row < nz * ny; row++) {
		const int z = z0 + row / ny, y = y0 + row % ny;
		for (int x = x0; x <= x1; x++)
			oref_rad_evaluate(h, x, y, z, out + ((size_t)row * nx + (x - x0)) * 12);
	}
}

/* GLRadiosityRenderer::EncodeValue (:465-515) */
uint32_t oref_rad_encode(void *p, const float v[3]) {
	return static_cast<Rad *>(p)->rad->EncodeValue(MakeVector3(v[0], v[1], v[2]));
}

/* chunk state: out = {dirty, minX, minY, minZ, maxX, maxY, maxZ} (GLRadiosityRenderer.h:52-59) */
void oref_rad_chunk_info(void *p, int cx, int cy, int cz, int out[7]) {
	auto &c = static_cast<Rad *>(p)->rad->GetChunk(cx, cy, cz);
	out[0] = c.dirty ? 1 : 0;
	out[1] = c.dirtyMinX;
	out[2] = c.dirtyMinY;
	out[3] = c.dirtyMinZ;
	out[4] = c.dirtyMaxX;
	out[5] = c.dirtyMaxY;
	out[6] = c.dirtyMaxZ;
}

int oref_rad_num_dirty(void *p) { return static_cast<Rad *>(p)->rad->GetNumDirtyChunks(); }

/* GLRadiosityRenderer::UpdateChunk (:517-543) */
void oref_rad_update_chunk(void *p, int cx, int cy, int cz) { static_cast<Rad *>(p)->rad->UpdateChunk(cx, cy, cz); }

/* UpdateChunk over every dirty chunk (the steady state the reference's 8-random-chunks-per-frame
   scheduler, :394-450, converges to); chunks are independent, so OpenMP over chunks.
   Returns the number of chunks updated. */
int oref_rad_update_all(void *p, int threads) {
	GLRadiosityRenderer *r = static_cast<Rad *>(p)->rad;
	const int n = (int)r->chunks.size();
	int cnt = 0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic, 1) reduction(+ : cnt)
	for (int i = 0; i < n; i++) {
		auto &c = r->chunks[i];
		if (c.dirty) {
			r->UpdateChunk(c.cx, c.cy, c.cz);
			cnt++;
		}
	}
	return cnt;
}

/* the four 16^3 RGB10A2 volumes of one chunk, [z][y][x] */
void oref_rad_read_chunk(void *p, int cx, int cy, int cz, uint32_t *flat, uint32_t *x, uint32_t *y, uint32_t *z) {
	auto &c = static_cast<Rad *>(p)->rad->GetChunk(cx, cy, cz);
	std::memcpy(flat, c.dataFlat, sizeof(c.dataFlat));
	std::memcpy(x, c.dataX, sizeof(c.dataX));
	std::memcpy(y, c.dataY, sizeof(c.dataY));
	std::memcpy(z, c.dataZ, sizeof(c.dataZ));
}

/* ---- GLAmbientShadowRenderer (Draw/GLAmbientShadowRenderer.cpp), CPU half: Evaluate :130-169, GameMapChanged /
   Invalidate :171-237, UpdateChunk :334-507.  Constructed over the same zeroed GLRenderer storage + NullGLDevice. */
void *oref_ao_create(void *map) {
	Ao *h = new Ao();
	std::memset(h->storage, 0, sizeof(h->storage));
	h->renderer = reinterpret_cast<GLRenderer *>(h->storage);
	h->map = static_cast<client::GameMap *>(map);
	new (&h->renderer->device) Handle<IGLDevice>(Handle<oref::NullGLDevice>::New().Cast<IGLDevice>());
	new (&h->renderer->settings) GLSettings();
	h->renderer->map = h->map;
	h->ao = new GLAmbientShadowRenderer(*h->renderer, *h->map);
	return h;
}

void oref_ao_free(void *p) {
	Ao *h = static_cast<Ao *>(p);
	delete h->ao;
	h->renderer->settings.~GLSettings();
	h->renderer->device.~Handle<IGLDevice>();
	delete h;
}

void oref_ao_get_rays(void *p, float out[48]) {
	auto &r = static_cast<Ao *>(p)->ao->rays;
	for (int i = 0; i < 16; i++) {
		out[i * 3] = r[i].x;
		out[i * 3 + 1] = r[i].y;
		out[i * 3 + 2] = r[i].z;
	}
}
void oref_ao_set_rays(void *p, const float in[48]) {
	auto &r = static_cast<Ao *>(p)->ao->rays;
	for (int i = 0; i < 16; i++)
		r[i] = MakeVector3(in[i * 3], in[i * 3 + 1], in[i * 3 + 2]);
}

float oref_ao_evaluate(void *p, int x, int y, int z) {
	return static_cast<Ao *>(p)->ao->Evaluate(IntVector3::Make(x, y, z));
}

void oref_ao_evaluate_box(void *p, int x0, int y0, int z0, int x1, int y1, int z1, float *out, int threads) {
	Ao *h = static_cast<Ao *>(p);
	const int nx = x1 - x0 + 1, ny = y1 - y0 + 1, nz = z1 - z0 + 1;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic, 4)
	for (int row = 0; row < nz * ny; row++) {
		const int z = z0 + row / ny, y = y0 + row % ny;
		for (int x = x0; x <= x1; x++)
			out[(size_t)row * nx + (x - x0)] = h->ao->Evaluate(IntVector3::Make(x, y, z));
	}
}

void oref_ao_map_changed(void *p, int x, int y, int z) {
	Ao *h = static_cast<Ao *>(p);
	h->ao->GameMapChanged(x, y, z, h->map);
}

void oref_ao_chunk_info(void *p, int cx, int cy, int cz, int out[7]) {
	auto &c = static_cast<Ao *>(p)->ao->GetChunk(cx, cy, cz);
	out[0] = c.dirty ? 1 : 0;
	out[1] = c.dirtyMinX;
	out[2] = c.dirtyMinY;
	out[3] = c.dirtyMinZ;
	out[4] = c.dirtyMaxX;
	out[5] = c.dirtyMaxY;
	out[6] = c.dirtyMaxZ;
}

int oref_ao_num_dirty(void *p) { return static_cast<Ao *>(p)->ao->GetNumDirtyChunks(); }

void oref_ao_update_chunk(void *p, int cx, int cy, int cz) { static_cast<Ao *>(p)->ao->UpdateChunk(cx, cy, cz); }

int oref_ao_update_all(void *p, int threads) {
	GLAmbientShadowRenderer *r = static_cast<Ao *>(p)->ao;
	const int n = (int)r->chunks.size();
	int cnt = 0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic, 1) reduction(+ : cnt)
	for (int i = 0; i < n; i++) {
		auto &c = r->chunks[i];
		if (c.dirty) {
			r->UpdateChunk(c.cx, c.cy, c.cz);
			cnt++;
		}
	}
	return cnt;
}

void oref_ao_read_chunk(void *p, int cx, int cy, int cz, float *out) {
	auto &c = static_cast<Ao *>(p)->ao->GetChunk(cx, cy, cz);
	std::memcpy(out, c.data, sizeof(c.data));
}

} // extern "C"
