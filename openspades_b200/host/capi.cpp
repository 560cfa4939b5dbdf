// C wrappers over the C++ host layer (VoxelMap, B200Renderer) so that the Python tests and bench.py
// can drive the same objects a C++ integrator would use.  Not part of the osb200 C ABI proper.
#include <cstring>
#include <exception>
#include <stdexcept>
#include <string>

#include "../../include/osb200.h"
#include "VoxelMap.h"

#define OSBH_API extern "C" __attribute__((visibility("default")))

namespace {
	thread_local std::string g_err;
	template <class F> int guarded(F f) {
		try {
			f();
			return 0;
		} catch (const std::exception &e) {
			g_err = e.what();
			return -1;
		}
	}
} // namespace

OSBH_API const char *osbh_last_error() { return g_err.c_str(); }

OSBH_API void *osbh_map_new() { return new osb200::VoxelMap(); }
OSBH_API void *osbh_map_load(const uint8_t *data, size_t len) {
	osb200::VoxelMap *m = nullptr;
	guarded([&] { m = osb200::VoxelMap::Load(data, len).release(); });
	return m;
}
OSBH_API void *osbh_map_load_file(const char *path) {
	osb200::VoxelMap *m = nullptr;
	guarded([&] { m = osb200::VoxelMap::LoadFile(path).release(); });
	return m;
}
OSBH_API void osbh_map_free(void *m) { delete static_cast<osb200::VoxelMap *>(m); }
OSBH_API const uint64_t *osbh_map_solid(void *m) { return static_cast<osb200::VoxelMap *>(m)->SolidData(); }
OSBH_API const uint32_t *osbh_map_color(void *m) { return static_cast<osb200::VoxelMap *>(m)->ColorData(); }
OSBH_API int osbh_map_set(void *m, int x, int y, int z, int solid, uint32_t color) {
	return guarded([&] { static_cast<osb200::VoxelMap *>(m)->Set(x, y, z, solid != 0, color); });
}
OSBH_API int osbh_map_set_arrays(void *m, const uint64_t *solid, const uint32_t *color) {
	auto *map = static_cast<osb200::VoxelMap *>(m);
	std::memcpy(map->MutableSolidData(), solid, sizeof(uint64_t) * 512 * 512);
	std::memcpy(map->MutableColorData(), color, sizeof(uint32_t) * 512 * 512 * 64);
	return 0;
}

// ---- B200Renderer ----
#include "B200Renderer.h"
#include "VoxelMapWrapper.h"

namespace {
	struct RendererBox {
		std::shared_ptr<osb200::MemoryPort> port;
		std::unique_ptr<osb200::B200Renderer> r;
	};
} // namespace

OSBH_API void *osbh_renderer_new(int w, int h, int device, int numerics, int swThreads, int undersampling,
                                 int radiosityMode) {
	RendererBox *b = nullptr;
	guarded([&] {
		auto box = std::unique_ptr<RendererBox>(new RendererBox());
		box->port = std::make_shared<osb200::MemoryPort>(w, h);
		osb200::B200RendererOptions o;
		o.device = device;
		o.numerics = numerics;
		o.swThreads = swThreads;
		o.undersampling = undersampling;
		o.radiosityMode = radiosityMode;
		box->r.reset(new osb200::B200Renderer(box->port, o));
		b = box.release();
	});
	return b;
}
OSBH_API void osbh_renderer_free(void *p) { delete static_cast<RendererBox *>(p); }
#define R(p) (static_cast<RendererBox *>(p)->r)
OSBH_API int osbh_renderer_init(void *p) { return guarded([&] { R(p)->Init(); }); }
OSBH_API int osbh_renderer_shutdown(void *p) { return guarded([&] { R(p)->Shutdown(); }); }
OSBH_API int osbh_renderer_set_map(void *p, void *map) {
	return guarded([&] { R(p)->SetGameMap(static_cast<osb200::VoxelMap *>(map)); });
}
// GameMap::Load on the device + SetGameMap; returns the new VoxelMap (owned by the caller) or NULL
OSBH_API void *osbh_renderer_load_map(void *p, const uint8_t *data, size_t len) {
	osb200::VoxelMap *m = nullptr;
	guarded([&] { m = R(p)->LoadGameMap(data, len).release(); });
	return m;
}
// GameMap::Save on the device: returns the stream length (< 0 on error), copies min(length, cap) bytes
OSBH_API long long osbh_renderer_save_map(void *p, uint8_t *out, size_t cap) {
	long long n = -1;
	guarded([&] {
		std::vector<uint8_t> v = R(p)->SaveGameMap();
		if (out && cap)
			std::memcpy(out, v.data(), v.size() < cap ? v.size() : cap);
		n = (long long)v.size();
	});
	return n;
}
OSBH_API int osbh_renderer_set_fog(void *p, float r, float g, float b, float dist) {
	return guarded([&] {
		R(p)->SetFogColor(osb200::Vector3{r, g, b});
		R(p)->SetFogDistance(dist);
	});
}
// scene = osb200_scene layout (origin, axis[3], fovX, fovY, zNear, zFar)
OSBH_API int osbh_renderer_start_scene(void *p, const osb200_scene *s) {
	return guarded([&] {
		osb200::SceneDefinition d;
		d.viewOrigin = osb200::Vector3{s->origin[0], s->origin[1], s->origin[2]};
		for (int i = 0; i < 3; i++)
			d.viewAxis[i] = osb200::Vector3{s->axis[i][0], s->axis[i][1], s->axis[i][2]};
		d.fovX = s->fovX;
		d.fovY = s->fovY;
		d.zNear = s->zNear;
		d.zFar = s->zFar;
		R(p)->StartScene(d);
	});
}
OSBH_API int osbh_renderer_add_sprite(void *p) {
	return guarded([&] { R(p)->AddSprite(nullptr, osb200::Vector3{0, 0, 0}, 1.f, 0.f); });
}
OSBH_API int osbh_renderer_end_scene(void *p) { return guarded([&] { R(p)->EndScene(); }); }
OSBH_API int osbh_renderer_frame_done(void *p) { return guarded([&] { R(p)->FrameDone(); }); }
OSBH_API int osbh_renderer_flip(void *p) { return guarded([&] { R(p)->Flip(); }); }
OSBH_API int osbh_renderer_swap_count(void *p) { return static_cast<RendererBox *>(p)->port->SwapCount(); }
OSBH_API int osbh_renderer_read_bitmap(void *p, uint32_t *out) {
	return guarded([&] {
		auto bm = R(p)->ReadBitmap();
		std::memcpy(out, bm->GetPixels(), (size_t)bm->GetWidth() * bm->GetHeight() * 4);
	});
}
OSBH_API int osbh_renderer_framebuffer(void *p, uint32_t *out) {
	return guarded([&] {
		auto &fb = static_cast<RendererBox *>(p)->port->GetFramebuffer();
		std::memcpy(out, fb.GetPixels(), (size_t)fb.GetWidth() * fb.GetHeight() * 4);
	});
}
OSBH_API int osbh_renderer_read_depth(void *p, float *out) {
	return guarded([&] {
		auto d = R(p)->ReadDepthBuffer();
		std::memcpy(out, d.data(), d.size() * 4);
	});
}
OSBH_API int osbh_renderer_screen(void *p, float *w, float *h) {
	return guarded([&] {
		*w = R(p)->ScreenWidth();
		*h = R(p)->ScreenHeight();
	});
}
OSBH_API int osbh_renderer_update_radiosity(void *p, int maxChunks, int *updated) {
	return guarded([&] { *updated = R(p)->UpdateRadiosity(maxChunks); });
}
OSBH_API int osbh_renderer_update_ambient(void *p, int maxChunks, int *updated) {
	return guarded([&] { *updated = R(p)->UpdateAmbient(maxChunks); });
}
OSBH_API int osbh_renderer_set_ambient_rays(void *p, const float *rays) {
	return guarded([&] { R(p)->SetAmbientRays(rays); });
}
OSBH_API int osbh_renderer_create_model(void *p, int w, int h, int d, const float *origin, const uint64_t *solid,
                                        const uint32_t *colors, void **out) {
	return guarded([&] {
		osb200::VoxelModelData vm;
		vm.width = w;
		vm.height = h;
		vm.depth = d;
		vm.origin = {origin[0], origin[1], origin[2]};
		vm.solidBits.assign(solid, solid + (size_t)w * h);
		vm.colors.assign(colors, colors + (size_t)w * h * d);
		*out = R(p)->CreateModel(vm).release();
	});
}
OSBH_API void osbh_model_free(void *m) { delete static_cast<osb200::Model *>(m); }
OSBH_API int osbh_renderer_render_model(void *p, void *model, const float *matrix, const float *customColor) {
	return guarded([&] {
		osb200::ModelRenderParam prm;
		for (int i = 0; i < 16; i++)
			prm.matrix[i] = matrix[i];
		prm.customColor = {customColor[0], customColor[1], customColor[2]};
		R(p)->RenderModel(*static_cast<osb200::Model *>(model), prm);
	});
}
OSBH_API int osbh_renderer_add_light(void *p, int type, const float *origin, float radius, const float *color) {
	return guarded([&] {
		osb200::DynamicLightParam prm;
		prm.type = static_cast<osb200::DynamicLightType>(type);
		prm.origin = {origin[0], origin[1], origin[2]};
		prm.radius = radius;
		prm.color = {color[0], color[1], color[2]};
		R(p)->AddLight(prm);
	});
}
OSBH_API int osbh_renderer_last_flushed(void *p) { return R(p)->LastFlushedColumns(); }
OSBH_API void *osbh_renderer_context(void *p) { return R(p)->Context(); }
OSBH_API int osbh_map_castray2(void *m, void *ctx, const float *v0, const float *dir, int maxSteps, osb200_rayhit *out,
                               size_t n) {
	return guarded([&] {
		static_cast<osb200::VoxelMap *>(m)->CastRay2Batch(static_cast<osb200_ctx *>(ctx), v0, dir, maxSteps, out, n);
	});
}
OSBH_API int osbh_map_castray(void *m, void *ctx, const float *v0, const float *dir, float length, int32_t *out, size_t n) {
	return guarded([&] {
		static_cast<osb200::VoxelMap *>(m)->CastRayBatch(static_cast<osb200_ctx *>(ctx), v0, dir, length, out, n);
	});
}
#undef R

// ---- VoxelMapWrapper (client::GameMapWrapper's surface) ----
OSBH_API void *osbh_wrapper_new(void *map, void *ctx) {
	return new osb200::VoxelMapWrapper(*static_cast<osb200::VoxelMap *>(map), static_cast<osb200_ctx *>(ctx));
}
OSBH_API void osbh_wrapper_free(void *w) { delete static_cast<osb200::VoxelMapWrapper *>(w); }
OSBH_API int osbh_wrapper_rebuild(void *w) {
	return guarded([&] { static_cast<osb200::VoxelMapWrapper *>(w)->Rebuild(); });
}
OSBH_API int osbh_wrapper_add_block(void *w, int x, int y, int z, uint32_t color) {
	return guarded([&] { static_cast<osb200::VoxelMapWrapper *>(w)->AddBlock(x, y, z, color); });
}
// World::ClusterizeBlocks through the wrapper: pieces written back to back into out (n cells), sizes[k] cells each;
// returns the number of pieces (< 0 on error, or when more than capPieces)
OSBH_API int osbh_wrapper_clusterize(void *w, const int16_t *cells, int n, int16_t *out, int32_t *sizes, int capPieces) {
	int count = -1;
	guarded([&] {
		std::vector<osb200::CellPos> in;
		for (int i = 0; i < n; i++)
			in.emplace_back(cells[i * 3], cells[i * 3 + 1], cells[i * 3 + 2]);
		auto pieces = static_cast<osb200::VoxelMapWrapper *>(w)->ClusterizeBlocks(in);
		if ((int)pieces.size() > capPieces)
			throw std::length_error("osbh_wrapper_clusterize: more pieces than capPieces");
		size_t o = 0;
		for (size_t k = 0; k < pieces.size(); k++) {
			sizes[k] = (int32_t)pieces[k].size();
			for (const auto &p : pieces[k]) {
				out[o * 3] = p.x;
				out[o * 3 + 1] = p.y;
				out[o * 3 + 2] = p.z;
				o++;
			}
		}
		count = (int)pieces.size();
	});
	return count;
}
// cells: n x {x, y, z} int16; returns the number of floating blocks (< 0 on error), writes min(count, cap)
OSBH_API int osbh_wrapper_remove_blocks(void *w, const int16_t *cells, int n, int16_t *out, int cap) {
	int count = -1;
	guarded([&] {
		std::vector<osb200::CellPos> in;
		for (int i = 0; i < n; i++)
			in.emplace_back(cells[i * 3], cells[i * 3 + 1], cells[i * 3 + 2]);
		std::vector<osb200::CellPos> fl = static_cast<osb200::VoxelMapWrapper *>(w)->RemoveBlocks(in);
		for (size_t i = 0; i < fl.size() && (int)i < cap; i++) {
			out[i * 3] = fl[i].x;
			out[i * 3 + 1] = fl[i].y;
			out[i * 3 + 2] = fl[i].z;
		}
		count = (int)fl.size();
	});
	return count;
}
