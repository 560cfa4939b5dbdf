/*
 * TEST INFRASTRUCTURE ONLY -- see oref.h.  Thin driver over the reference's own classes,
 * following the head-less pattern of Client/HitTestDebugger.cpp:38-57,147-305.
 */
#include "oref.h"

#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include <Client/GameMap.h>
#include <Client/GameMapWrapper.h>
#include <Client/SceneDefinition.h>
#include <Core/Bitmap.h>
#include <Core/Exception.h>
#include <Core/MemoryStream.h>
#include <Core/DynamicMemoryStream.h>
#include <Core/Settings.h>
#include <Core/StdStream.h>
#include <Draw/SWFeatureLevel.h>
#include <Core/VoxelModel.h>
#include <Draw/SWFlatMapRenderer.h>
#include <Draw/SWImage.h>
#include <Draw/SWMapRenderer.h>
#include <Draw/SWModel.h>
#include <Draw/SWPort.h>
#include <Draw/SWRenderer.h>

using namespace spades;

/* SWMapRenderer::Line / LinePixel are private to SWMapRenderer.cpp (:70-105); the harness only
   needs to read them, so it names layout-identical definitions here (8-byte pixel: combined,
   depth; Line: vector, horizonDir, pitchTanMin, pitchScale, pitchTanMinI, pitchScaleI). */
struct spades::draw::SWMapRenderer::LinePixel {
	uint32_t combined;
	float depth;
};
struct spades::draw::SWMapRenderer::Line {
	std::vector<LinePixel> pixels;
	Vector3 horizonDir;
	float pitchTanMin;
	float pitchScale;
	int pitchTanMinI;
	int pitchScaleI;
};

namespace {
	thread_local std::string g_err;

	class Port : public draw::SWPort {
		Handle<Bitmap> bmp;

	public:
		Port(int w, int h) { bmp = Handle<Bitmap>::New(w, h); }
		Bitmap &GetFramebuffer() override { return *bmp; }
		void Swap() override {}
	};

	struct Renderer {
		Handle<Port> port;
		Handle<draw::SWRenderer> r;
		int w, h;
	};
} // namespace

extern "C" {

const char *oref_variant(void) {
#if ENABLE_SSE2
	return "sse2";
#else
	return "ieee";
#endif
}

const char *oref_last_error(void) { return g_err.c_str(); }

void *oref_map_load_vxl(const char *path) {
	try {
		FILE *f = fopen(path, "rb");
		if (!f) { g_err = std::string("cannot open ") + path; return nullptr; }
		StdStream s(f, true);
		return client::GameMap::Load(&s);
	} catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

void *oref_map_load_vxl_mem(const void *data, size_t len) {
	try {
		MemoryStream s(reinterpret_cast<const char *>(data), len);
		return client::GameMap::Load(&s);
	} catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

long long oref_map_save_vxl(void *map, void *out, size_t cap) {
	try {
		DynamicMemoryStream s;
		reinterpret_cast<client::GameMap *>(map)->Save(&s);
		const uint64_t n = s.GetLength();
		s.SetPosition(0);
		if (out && cap)
			s.Read(out, (size_t)(n < cap ? n : cap));
		return (long long)n;
	} catch (const std::exception &e) { g_err = e.what(); return -1; }
}

void *oref_map_from_arrays(const uint64_t *solid, const uint32_t *color) {
	auto *m = new client::GameMap();
	std::memcpy(m->solidMap, solid, sizeof(m->solidMap));
	std::memcpy(m->colorMap, color, sizeof(m->colorMap));
	return m;
}

void oref_map_free(void *map) { reinterpret_cast<client::GameMap *>(map)->Release(); }

void oref_map_get_arrays(void *map, uint64_t *solid, uint32_t *color) {
	auto *m = reinterpret_cast<client::GameMap *>(map);
	if (solid) std::memcpy(solid, m->solidMap, sizeof(m->solidMap));
	if (color) std::memcpy(color, m->colorMap, sizeof(m->colorMap));
}

void oref_map_set(void *map, int x, int y, int z, int solid, uint32_t color) {
	reinterpret_cast<client::GameMap *>(map)->Set(x, y, z, solid != 0, color);
}

void oref_castray2(void *map, const float *v0, const float *dir, int maxSteps, oref_rayhit *out,
                   size_t n, int threads) {
	auto *m = reinterpret_cast<const client::GameMap *>(map);
	if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static)
	for (long long i = 0; i < (long long)n; i++) {
		client::GameMap::RayCastResult res;
		std::memset(&res, 0, sizeof(res)); /* hitBlock/normal stay 0 when maxSteps==0 */
		res = m->CastRay2(MakeVector3(v0[i * 3], v0[i * 3 + 1], v0[i * 3 + 2]),
		                  MakeVector3(dir[i * 3], dir[i * 3 + 1], dir[i * 3 + 2]), maxSteps);
		oref_rayhit &o = out[i];
		o.hit = res.hit; o.startSolid = res.startSolid; o.pad[0] = o.pad[1] = 0;
		o.hitPos[0] = res.hitPos.x; o.hitPos[1] = res.hitPos.y; o.hitPos[2] = res.hitPos.z;
		o.hitBlock[0] = res.hitBlock.x; o.hitBlock[1] = res.hitBlock.y; o.hitBlock[2] = res.hitBlock.z;
		o.normal[0] = res.normal.x; o.normal[1] = res.normal.y; o.normal[2] = res.normal.z;
	}
}

/* ---- GameMapWrapper (Client/GameMapWrapper.cpp): floating-block detection ---- */
void *oref_wrapper_create(void *map) {
	try {
		auto *w = new client::GameMapWrapper(*reinterpret_cast<client::GameMap *>(map));
		w->Rebuild();
		return w;
	} catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}
void oref_wrapper_free(void *w) { delete reinterpret_cast<client::GameMapWrapper *>(w); }
void oref_wrapper_add_block(void *w, int x, int y, int z, uint32_t color) {
	reinterpret_cast<client::GameMapWrapper *>(w)->AddBlock(x, y, z, color);
}
/* cells: n x {x, y, z} int16; returns the number of floating blocks, writes min(count, cap) of them */
int oref_wrapper_remove_blocks(void *w, const int16_t *cells, int n, int16_t *out, int cap) {
	try {
		std::vector<client::CellPos> in;
		for (int i = 0; i < n; i++)
			in.emplace_back(cells[i * 3], cells[i * 3 + 1], cells[i * 3 + 2]);
		std::vector<client::CellPos> fl = reinterpret_cast<client::GameMapWrapper *>(w)->RemoveBlocks(in);
		for (size_t i = 0; i < fl.size() && (int)i < cap; i++) {
			out[i * 3] = fl[i].x;
			out[i * 3 + 1] = fl[i].y;
			out[i * 3 + 2] = fl[i].z;
		}
		return (int)fl.size();
	} catch (const std::exception &e) { g_err = e.what(); return -1; }
}
/* linkMap != Invalid per voxel, as bit z of out[x][y] */
void oref_wrapper_linked(void *w, uint64_t *out) {
	auto *gw = reinterpret_cast<client::GameMapWrapper *>(w);
	for (int x = 0; x < 512; x++)
		for (int y = 0; y < 512; y++) {
			uint64_t b = 0;
			for (int z = 0; z < 64; z++)
				if (gw->linkMap[(x * 512 + y) * 64 + z] != 0)
					b |= 1ull << z;
			out[x * 512 + y] = b;
		}
}

/* GameMap::CastRay (GameMap.cpp:208-359) */
void oref_castray(void *map, const float *v0, const float *dir, float length, int32_t *out, size_t n, int threads) {
	auto *m = reinterpret_cast<const client::GameMap *>(map);
	if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static)
	for (long long i = 0; i < (long long)n; i++) {
		IntVector3 hb = IntVector3::Make(0, 0, 0);
		const bool hit = m->CastRay(MakeVector3(v0[i * 3], v0[i * 3 + 1], v0[i * 3 + 2]),
		                            MakeVector3(dir[i * 3], dir[i * 3 + 1], dir[i * 3 + 2]), length, hb);
		out[i * 4] = hit ? 1 : 0;
		out[i * 4 + 1] = hit ? hb.x : 0;
		out[i * 4 + 2] = hit ? hb.y : 0;
		out[i * 4 + 3] = hit ? hb.z : 0;
	}
}

void oref_renderer_set_threads(int threads) {
	Settings::ItemHandle h("r_swNumThreads", nullptr);
	h = threads;
}

void oref_renderer_set_undersampling(int under) { /* cvar r_swUndersampling (SWMapRenderer.cpp:37), read on every frame */
	Settings::ItemHandle h("r_swUndersampling", nullptr);
	h = under;
}

void *oref_renderer_create(int w, int h, int level, int threads) {
	try {
		oref_renderer_set_threads(threads);
		auto *R = new Renderer();
		R->w = w; R->h = h;
		R->port = Handle<Port>::New(w, h);
		if (level < 0)
			R->r = Handle<draw::SWRenderer>::New(R->port.Cast<draw::SWPort>());
		else
			R->r = Handle<draw::SWRenderer>::New(R->port.Cast<draw::SWPort>(),
			                                     static_cast<draw::SWFeatureLevel>(level));
		R->r->Init();
		return R;
	} catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

int oref_renderer_level(void *r) {
	return static_cast<int>(reinterpret_cast<Renderer *>(r)->r->featureLevel);
}

void oref_renderer_free(void *r) {
	auto *R = reinterpret_cast<Renderer *>(r);
	try { R->r->SetGameMap({}); R->r->Shutdown(); } catch (...) {}
	delete R;
}

void oref_renderer_set_map(void *r, void *map) {
	auto *R = reinterpret_cast<Renderer *>(r);
	if (map) R->r->SetGameMap(*reinterpret_cast<client::GameMap *>(map));
	else R->r->SetGameMap({});
}

/* SWFlatMapRenderer's image (GeneratePixel per column, SWFlatMapRenderer.cpp:52-105), after draining pending updates */
int oref_renderer_minimap(void *r, uint32_t *out /*[y][x] 512x512*/) {
	auto *R = reinterpret_cast<Renderer *>(r);
	if (!R->r->flatMapRenderer) { g_err = "no map set"; return -1; }
	R->r->flatMapRenderer->Update();
	std::memcpy(out, R->r->flatMapRenderer->img->GetRawBitmap(), sizeof(uint32_t) * 512 * 512);
	return 0;
}

void oref_renderer_set_fog(void *r, float cr, float cg, float cb, float dist) {
	auto *R = reinterpret_cast<Renderer *>(r);
	R->r->SetFogColor(MakeVector3(cr, cg, cb));
	R->r->SetFogDistance(dist);
}

int oref_renderer_render(void *r, const oref_scene *s, uint32_t *bgrx, float *depth) {
	auto *R = reinterpret_cast<Renderer *>(r);
	try {
		client::SceneDefinition def;
		def.viewportLeft = def.viewportTop = 0;
		def.viewportWidth = R->w; def.viewportHeight = R->h;
		def.fovX = s->fovX; def.fovY = s->fovY;
		def.viewOrigin = MakeVector3(s->origin[0], s->origin[1], s->origin[2]);
		for (int i = 0; i < 3; i++)
			def.viewAxis[i] = MakeVector3(s->axis[i][0], s->axis[i][1], s->axis[i][2]);
		def.zNear = s->zNear; def.zFar = s->zFar;
		def.skipWorld = false;
		R->r->StartScene(def);
		R->r->EndScene();
		R->r->FrameDone();
		size_t n = (size_t)R->w * R->h;
		if (bgrx) std::memcpy(bgrx, R->port->GetFramebuffer().GetPixels(), n * 4);
		if (depth) std::memcpy(depth, R->r->depthBuffer.data(), n * 4);
		return 0;
	} catch (const std::exception &e) { g_err = e.what(); return -1; }
}

int oref_renderer_rle(void *r, int x, int y, int8_t *out, int cap) {
	auto *R = reinterpret_cast<Renderer *>(r);
	auto *mr = R->r->mapRenderer.get();
	if (!mr) return -1;
	int idx = x + y * mr->w;
	int len = (int)mr->rleLen[idx];
	const int8_t *p = mr->rleHeap.Dereference<int8_t>(mr->rle[idx]);
	if (len > cap) len = cap;
	std::memcpy(out, p, len);
	return len;
}

int oref_renderer_lines(void *r, int *lineRes) {
	auto *R = reinterpret_cast<Renderer *>(r);
	auto *mr = R->r->mapRenderer.get();
	if (!mr) return -1;
	if (lineRes) *lineRes = (int)mr->lineResolution;
	return (int)mr->lines.size();
}

int oref_renderer_line(void *r, int i, uint32_t *combined, float *depth, float *misc, int32_t *misci) {
	auto *R = reinterpret_cast<Renderer *>(r);
	auto *mr = R->r->mapRenderer.get();
	if (!mr || i < 0 || i >= (int)mr->lines.size()) return -1;
	auto &l = mr->lines[i];
	int n = (int)l.pixels.size();
	for (int j = 0; j < n; j++) {
		if (combined) combined[j] = l.pixels[j].combined;
		if (depth) depth[j] = l.pixels[j].depth;
	}
	if (misc) { misc[0] = l.horizonDir.x; misc[1] = l.horizonDir.y; misc[2] = l.pitchTanMin; misc[3] = l.pitchScale; }
	if (misci) { misci[0] = l.pitchTanMinI; misci[1] = l.pitchScaleI; }
	return n;
}

/* ---- voxel models + dynamic lights ---- */
void *oref_vmodel_load_kv6(const void *data, size_t len) {
	try {
		MemoryStream ms(reinterpret_cast<char *>(const_cast<void *>(data)), len);
		Handle<VoxelModel> vm = VoxelModel::LoadKV6(ms);
		return std::move(vm).Unmanage();
	} catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

void *oref_vmodel_from_arrays(int w, int h, int d, const float origin[3], const uint64_t *solid, const uint32_t *colors) {
	try {
		Handle<VoxelModel> vm = Handle<VoxelModel>::New(w, h, d);
		vm->SetOrigin(MakeVector3(origin[0], origin[1], origin[2]));
		for (int y = 0; y < h; y++)
			for (int x = 0; x < w; x++) {
				vm->GetSolidBitsAt(x, y) = solid[x + y * w];
				for (int z = 0; z < d; z++)
					vm->GetColor(x, y, z) = colors[(size_t)(x + y * w) * d + z];
			}
		return std::move(vm).Unmanage();
	} catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

void oref_vmodel_free(void *vm) { reinterpret_cast<VoxelModel *>(vm)->Release(); }

void oref_vmodel_info(void *p, int dims[3], float origin[3]) {
	auto *vm = reinterpret_cast<VoxelModel *>(p);
	dims[0] = vm->GetWidth(); dims[1] = vm->GetHeight(); dims[2] = vm->GetDepth();
	Vector3 o = vm->GetOrigin();
	origin[0] = o.x; origin[1] = o.y; origin[2] = o.z;
}

void oref_vmodel_arrays(void *p, uint64_t *solid, uint32_t *colors) {
	auto *vm = reinterpret_cast<VoxelModel *>(p);
	const int w = vm->GetWidth(), h = vm->GetHeight(), d = vm->GetDepth();
	for (int y = 0; y < h; y++)
		for (int x = 0; x < w; x++) {
			solid[x + y * w] = vm->GetSolidBitsAt(x, y);
			for (int z = 0; z < d; z++)
				colors[(size_t)(x + y * w) * d + z] = vm->GetColor(x, y, z);
		}
}

void *oref_model_create(void *r, void *vm) {
	auto *R = reinterpret_cast<Renderer *>(r);
	try {
		Handle<client::IModel> m = R->r->CreateModel(*reinterpret_cast<VoxelModel *>(vm));
		return std::move(m).Unmanage();
	} catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

void oref_model_free(void *m) { reinterpret_cast<client::IModel *>(m)->Release(); }

int oref_model_render_data(void *m, uint32_t *data, int dataCap, uint32_t *addr, int addrCap) {
	auto *sm = dynamic_cast<draw::SWModel *>(reinterpret_cast<client::IModel *>(m));
	if (!sm) return -1;
	if (data && dataCap >= (int)sm->renderData.size())
		std::memcpy(data, sm->renderData.data(), sm->renderData.size() * 4);
	if (addr && addrCap >= (int)sm->renderDataAddr.size())
		std::memcpy(addr, sm->renderDataAddr.data(), sm->renderDataAddr.size() * 4);
	return (int)sm->renderData.size();
}

int oref_renderer_render_scene(void *r, const oref_scene *s, void *const *models, const oref_model_param *params,
                               int nModels, const oref_light *lights, int nLights, uint32_t *bgrx, float *depth) {
	auto *R = reinterpret_cast<Renderer *>(r);
	try {
		client::SceneDefinition def;
		def.viewportLeft = def.viewportTop = 0;
		def.viewportWidth = R->w; def.viewportHeight = R->h;
		def.fovX = s->fovX; def.fovY = s->fovY;
		def.viewOrigin = MakeVector3(s->origin[0], s->origin[1], s->origin[2]);
		for (int i = 0; i < 3; i++)
			def.viewAxis[i] = MakeVector3(s->axis[i][0], s->axis[i][1], s->axis[i][2]);
		def.zNear = s->zNear; def.zFar = s->zFar;
		def.skipWorld = false;
		R->r->StartScene(def);
		for (int i = 0; i < nModels; i++) {
			client::ModelRenderParam p;
			for (int k = 0; k < 16; k++) p.matrix.m[k] = params[i].matrix[k];
			p.customColor = MakeVector3(params[i].customColor[0], params[i].customColor[1], params[i].customColor[2]);
			R->r->RenderModel(*reinterpret_cast<client::IModel *>(models[i]), p);
		}
		for (int i = 0; i < nLights; i++) {
			client::DynamicLightParam p;
			p.type = client::DynamicLightTypePoint;
			p.origin = MakeVector3(lights[i].origin[0], lights[i].origin[1], lights[i].origin[2]);
			p.radius = lights[i].radius;
			p.color = MakeVector3(lights[i].color[0], lights[i].color[1], lights[i].color[2]);
			R->r->AddLight(p);
		}
		R->r->EndScene();
		R->r->FrameDone();
		size_t n = (size_t)R->w * R->h;
		if (bgrx) std::memcpy(bgrx, R->port->GetFramebuffer().GetPixels(), n * 4);
		if (depth) std::memcpy(depth, R->r->depthBuffer.data(), n * 4);
		return 0;
	} catch (const std::exception &e) { g_err = e.what(); return -1; }
}

} // extern "C"
