// See B200Renderer.h.  Mirrors draw::SWRenderer's state machine (Sources/Draw/SWRenderer.cpp) method by method; every
// per-pixel / per-voxel computation is an osb200_* call into the CUDA library -- there is no CPU rendering path here.
#include "B200Renderer.h"

#include <cstring>

#include <Client/GameMap.h>
#include <Core/Bitmap.h>
#include <Core/Debug.h>
#include <Core/Exception.h>
#include <Core/VoxelModel.h>
#include <Core/VoxelModelLoader.h>
#include <Draw/SWFeatureLevel.h>

namespace spades {
	namespace draw {

		B200Model::~B200Model() { osb200_model_destroy(model); }

		int B200Renderer::NumericsOfThisBuild() {
			const int level = static_cast<int>(DetectFeatureLevel());
#if ENABLE_SSE2
			if (level >= static_cast<int>(SWFeatureLevel::SSE2))
				return OSB200_NUM_SSE2;
#endif
#if ENABLE_SSE
			if (level >= static_cast<int>(SWFeatureLevel::SSE))
				return OSB200_NUM_SSE;
#endif
			(void)level;
			return OSB200_NUM_IEEE;
		}

		B200Renderer::B200Renderer(Handle<SWPort> port, int numerics, int device) : port(port) {
			SPADES_MARK_FUNCTION();
			if (!port)
				SPRaise("Port is null."); // SWRenderer.cpp:61
			osb200_opts opts;
			std::memset(&opts, 0, sizeof(opts));
			opts.device = device;
			opts.numerics = numerics;
			opts.sw_threads = 0;    // r_swNumThreads' default (4, SWRenderer.cpp:41)
			opts.undersampling = 0; // r_swUndersampling's default (off, SWMapRenderer.cpp:37)
			if (osb200_create(&opts, &ctx) != 0)
				SPRaise("B200Renderer: %s", osb200_last_error());
			SetFramebuffer(&port->GetFramebuffer());
		}

		B200Renderer::~B200Renderer() {
			SPADES_MARK_FUNCTION();
			Shutdown();
			models.clear();
			osb200_destroy(ctx);
			ctx = nullptr;
		}

		void B200Renderer::RaiseAbi(const char *what) { SPRaise("%s: %s", what, osb200_last_error()); }

		void B200Renderer::SetFramebuffer(Bitmap *bmp) {
			// SWRenderer.cpp:95-108
			if (bmp == nullptr)
				SPRaise("Framebuffer is null.");
			if (fb) {
				SPAssert(bmp->GetWidth() == fb->GetWidth());
				SPAssert(bmp->GetHeight() == fb->GetHeight());
			}
			fb = bmp;
			if ((bmp->GetWidth() & 7) || (bmp->GetHeight() & 7))
				SPRaise("Framebuffer size is not multiple of 8.");
		}

		void B200Renderer::Init() {
			SPADES_MARK_FUNCTION();
			inited = true;
		}

		void B200Renderer::Shutdown() {
			SPADES_MARK_FUNCTION();
			SetGameMap(nullptr);
			map = nullptr;
			port = nullptr;
			inited = false;
		}

		void B200Renderer::EnsureSceneStarted() {
			if (!duringSceneRendering)
				SPRaise("Illegal call outside of StartScene ... EndScene");
		}
		void B200Renderer::EnsureSceneNotStarted() {
			if (duringSceneRendering)
				SPRaise("Illegal call between StartScene ... EndScene");
		}
		void B200Renderer::EnsureInitialized() {
			if (!inited)
				SPRaise("Renderer is not initialized");
			EnsureValid();
		}
		void B200Renderer::EnsureValid() {
			if (!port)
				SPRaise("Renderer is not valid");
		}

		Handle<client::IImage> B200Renderer::RegisterImage(const char *) {
			SPADES_MARK_FUNCTION();
			EnsureValid();
			return Handle<B200Image>::New(1.f, 1.f).Cast<client::IImage>();
		}
		Handle<client::IImage> B200Renderer::CreateImage(Bitmap &bmp) {
			SPADES_MARK_FUNCTION();
			EnsureValid();
			return Handle<B200Image>::New((float)bmp.GetWidth(), (float)bmp.GetHeight()).Cast<client::IImage>();
		}

		Handle<client::IModel> B200Renderer::RegisterModel(const char *filename) {
			SPADES_MARK_FUNCTION();
			EnsureInitialized();
			auto vm = VoxelModelLoader::Load(filename); // SWModelManager::RegisterModel (SWModel.cpp:128-141), uncached
			return CreateModel(*vm);
		}

		Handle<client::IModel> B200Renderer::CreateModel(VoxelModel &vm) {
			SPADES_MARK_FUNCTION();
			EnsureInitialized();
			// SWModel's constructor (SWModel.cpp:25-100) runs inside osb200_model_create; hand it the voxel data in the
			// order that constructor reads it: solid[x + y*w] bit z, colors[(x + y*w)*d + z]
			const int w = vm.GetWidth(), h = vm.GetHeight(), d = vm.GetDepth();
			std::vector<uint64_t> solid((size_t)w * h);
			std::vector<uint32_t> colors((size_t)w * h * d, 0u);
			for (int y = 0; y < h; y++)
				for (int x = 0; x < w; x++) {
					const uint64_t bits = vm.GetSolidBitsAt(x, y);
					solid[x + (size_t)y * w] = bits;
					for (int z = 0; z < d; z++)
						if ((bits >> z) & 1ull)
							colors[(x + (size_t)y * w) * d + z] = vm.GetColor(x, y, z);
				}
			const Vector3 o = vm.GetOrigin();
			const float origin[3] = {o.x, o.y, o.z};
			osb200_model *m = nullptr;
			if (osb200_model_create(ctx, w, h, d, origin, solid.data(), colors.data(), &m) != 0)
				RaiseAbi("osb200_model_create");
			// SWModel's bounding box (SWModel.cpp:41-47,108-113): the voxel grid shifted by origin - 0.5
			AABB3 box;
			box.min = MakeVector3(o.x - .5f, o.y - .5f, o.z - .5f);
			box.max = MakeVector3((float)w + o.x - .5f, (float)h + o.y - .5f, (float)d + o.z - .5f);
			return Handle<B200Model>::New(m, box).Cast<client::IModel>();
		}

		void B200Renderer::SetGameMap(stmp::optional<client::GameMap &> newMap) {
			SPADES_MARK_FUNCTION();
			// SWRenderer.cpp:172-191
			if (newMap)
				EnsureInitialized();
			if (newMap.get_pointer() == map.GetPointerOrNull())
				return;
			if (map)
				map->RemoveListener(this);
			map = newMap;
			if (!map) {
				if (ctx)
					osb200_map_clear(ctx);
				return;
			}
			SPAssert(map->Width() == OSB200_MAP_W && map->Height() == OSB200_MAP_H && map->Depth() == OSB200_MAP_D);
			map->AddListener(this);
			// SWMapRenderer's constructor (RLE build, SWMapRenderer.cpp:107-138) becomes: mirror + upload of the whole map
			const int W = OSB200_MAP_W, H = OSB200_MAP_H, D = OSB200_MAP_D;
			solidMirror.resize((size_t)W * H);
			colorMirror.resize((size_t)W * H * D);
			for (int x = 0; x < W; x++)
				for (int y = 0; y < H; y++) {
					solidMirror[(size_t)x * H + y] = map->GetSolidMapWrapped(x, y);
					uint32_t *col = &colorMirror[((size_t)x * H + y) * D];
					for (int z = 0; z < D; z++)
						col[z] = map->GetColor(x, y, z);
				}
			if (osb200_map_upload(ctx, solidMirror.data(), colorMirror.data()) != 0)
				RaiseAbi("osb200_map_upload");
		}

		void B200Renderer::GameMapChanged(int x, int y, int z, client::GameMap *m) {
			// SWRenderer.cpp:1148-1154 -> SWFlatMapRenderer::SetNeedsUpdate: O(1) under a mutex.  Called from GameMap::Set
			// after the voxel changed: refresh the mirror's column word and voxel colour, mark the column dirty.
			if (m != map.GetPointerOrNull())
				return;
			std::lock_guard<std::mutex> guard(dirtyMutex);
			const size_t c = (size_t)x * OSB200_MAP_H + y;
			solidMirror[c] = m->GetSolidMapWrapped(x, y);
			colorMirror[c * OSB200_MAP_D + z] = m->GetColor(x, y, z);
			osb200_map_mark_dirty(ctx, x, y, z);
		}

		void B200Renderer::SetFogColor(Vector3 v) { fogColor = v; }

		void B200Renderer::StartScene(const client::SceneDefinition &def) {
			SPADES_MARK_FUNCTION();
			EnsureInitialized();
			EnsureSceneNotStarted();
			sceneDef = def;
			duringSceneRendering = true;
		}

		void B200Renderer::RenderModel(client::IModel &model, const client::ModelRenderParam &param) {
			SPADES_MARK_FUNCTION();
			// SWRenderer.cpp:565-576
			EnsureInitialized();
			EnsureSceneStarted();
			B200Model *m = dynamic_cast<B200Model *>(&model);
			if (!m)
				SPInvalidArgument("model");
			QueuedModel q;
			q.model = Handle<B200Model>(m, true);
			q.param = param;
			models.push_back(q);
		}

		void B200Renderer::AddLight(const client::DynamicLightParam &param) {
			SPADES_MARK_FUNCTION();
			// SWRenderer.cpp:578-585: only point lights; the culls and the screen rectangle (:587-649) run inside
			// osb200_render_scene with the same arithmetic
			EnsureInitialized();
			EnsureSceneStarted();
			if (param.type != client::DynamicLightTypePoint)
				return;
			osb200_light l;
			l.origin[0] = param.origin.x;
			l.origin[1] = param.origin.y;
			l.origin[2] = param.origin.z;
			l.radius = param.radius;
			l.color[0] = param.color.x;
			l.color[1] = param.color.y;
			l.color[2] = param.color.z;
			lights.push_back(l);
		}

		// sprites, debug lines and the 2-D image calls keep SWRenderer's call-order rules and draw nothing
		void B200Renderer::AddDebugLine(Vector3, Vector3, Vector4) {
			EnsureInitialized();
			EnsureSceneStarted();
		}
		void B200Renderer::AddSprite(client::IImage &, Vector3, float, float) {
			EnsureInitialized();
			EnsureSceneStarted();
		}
		void B200Renderer::AddLongSprite(client::IImage &, Vector3, Vector3, float) {
			EnsureInitialized();
			EnsureSceneStarted();
		}
		void B200Renderer::MultiplyScreenColor(Vector3) {
			EnsureInitialized();
			EnsureSceneNotStarted();
		}
		void B200Renderer::SetColor(Vector4) { EnsureValid(); }
		void B200Renderer::SetColorAlphaPremultiplied(Vector4) { EnsureValid(); }
		void B200Renderer::DrawImage(stmp::optional<client::IImage &>, const Vector2 &) {
			EnsureValid();
			EnsureSceneNotStarted();
		}
		void B200Renderer::DrawImage(stmp::optional<client::IImage &>, const AABB2 &) {
			EnsureValid();
			EnsureSceneNotStarted();
		}
		void B200Renderer::DrawImage(stmp::optional<client::IImage &>, const Vector2 &, const AABB2 &) {
			EnsureValid();
			EnsureSceneNotStarted();
		}
		void B200Renderer::DrawImage(stmp::optional<client::IImage &>, const AABB2 &, const AABB2 &) {
			EnsureValid();
			EnsureSceneNotStarted();
		}
		void B200Renderer::DrawImage(stmp::optional<client::IImage &>, const Vector2 &, const Vector2 &, const Vector2 &,
		                             const AABB2 &) {
			EnsureValid();
			EnsureSceneNotStarted();
		}
		void B200Renderer::DrawFlatGameMap(const AABB2 &, const AABB2 &) {
			EnsureValid();
			EnsureSceneNotStarted();
		}

		void B200Renderer::EndScene() {
			SPADES_MARK_FUNCTION();
			// SWRenderer.cpp:709-745: clear, SWFlatMapRenderer::Update (dirty columns -> UpdateRle), terrain, models,
			// dynamic lights, fog.  The clear is folded into the kernels (the terrain pass writes every pixel, or 0x7f7f7f
			// when the eye is inside a solid voxel / no map is set, SWMapRenderer.cpp:1181-1184).
			EnsureInitialized();
			EnsureSceneStarted();
			duringSceneRendering = false;
			if (map) {
				{
					std::lock_guard<std::mutex> guard(dirtyMutex);
					lastFlushed = osb200_map_flush(ctx, solidMirror.data(), colorMirror.data());
				}
				if (lastFlushed < 0)
					RaiseAbi("osb200_map_flush");
			}
			osb200_scene sc;
			sc.origin[0] = sceneDef.viewOrigin.x;
			sc.origin[1] = sceneDef.viewOrigin.y;
			sc.origin[2] = sceneDef.viewOrigin.z;
			for (int i = 0; i < 3; i++) {
				sc.axis[i][0] = sceneDef.viewAxis[i].x;
				sc.axis[i][1] = sceneDef.viewAxis[i].y;
				sc.axis[i][2] = sceneDef.viewAxis[i].z;
			}
			sc.fovX = sceneDef.fovX;
			sc.fovY = sceneDef.fovY;
			sc.zNear = sceneDef.zNear;
			sc.zFar = sceneDef.zFar;
			std::vector<osb200_model_instance> inst(models.size());
			for (size_t i = 0; i < models.size(); i++) {
				inst[i].model = models[i].model->model;
				std::memcpy(inst[i].matrix, models[i].param.matrix.m, sizeof(float) * 16);
				inst[i].customColor[0] = models[i].param.customColor.x;
				inst[i].customColor[1] = models[i].param.customColor.y;
				inst[i].customColor[2] = models[i].param.customColor.z;
			}
			const float fog[3] = {fogColor.x, fogColor.y, fogColor.z};
			const int rc = osb200_render_scene(ctx, &sc, fb->GetWidth(), fb->GetHeight(), fog, fogDistance, inst.data(),
			                                   (int)inst.size(), lights.data(), (int)lights.size(), fb->GetPixels(), nullptr);
			models.clear(); // :731,737 -- whether or not the draw succeeded
			lights.clear();
			if (rc != 0)
				RaiseAbi("osb200_render_scene");
		}

		void B200Renderer::FrameDone() {
			SPADES_MARK_FUNCTION();
			EnsureValid();
			EnsureSceneNotStarted();
		}

		void B200Renderer::Flip() {
			SPADES_MARK_FUNCTION();
			// SWRenderer.cpp:1045-1076: the pixel pointer may change every frame
			EnsureValid();
			EnsureSceneNotStarted();
			port->Swap();
			SetFramebuffer(&port->GetFramebuffer());
		}

		Handle<Bitmap> B200Renderer::ReadBitmap() {
			SPADES_MARK_FUNCTION();
			// SWRenderer.cpp:1078-1098: RGBA, bottom-up, alpha forced to 0xff
			EnsureValid();
			EnsureSceneNotStarted();
			const int w = fb->GetWidth(), h = fb->GetHeight();
			uint32_t *inPix = fb->GetPixels();
			auto bm = Handle<Bitmap>::New(w, h);
			uint32_t *outPix = bm->GetPixels();
			for (int y = 0; y < h; y++) {
				const uint32_t *src = inPix + (size_t)y * w;
				uint32_t *dest = outPix + (size_t)(h - 1 - y) * w;
				for (int x = 0; x < w; x++) {
					const uint32_t c = src[x];
					dest[x] = 0xff000000u | (c & 0xff00u) | ((c & 0xffu) << 16) | ((c & 0xff0000u) >> 16);
				}
			}
			return bm;
		}

		float B200Renderer::ScreenWidth() {
			EnsureValid();
			return static_cast<float>(fb->GetWidth());
		}
		float B200Renderer::ScreenHeight() {
			EnsureValid();
			return static_cast<float>(fb->GetHeight());
		}

		std::vector<float> B200Renderer::ReadDepthBuffer() {
			EnsureValid();
			std::vector<float> d((size_t)fb->GetWidth() * fb->GetHeight(), 0.f);
			if (osb200_read_depth(ctx, d.data(), d.size()) != 0)
				RaiseAbi("osb200_read_depth");
			return d;
		}

		void B200Renderer::CastRay2Batch(const Vector3 *v0, const Vector3 *dir, int maxSteps,
		                                 client::GameMap::RayCastResult *out, size_t n) {
			static_assert(sizeof(Vector3) == 12, "Vector3 is three packed floats");
			static_assert(sizeof(client::GameMap::RayCastResult) == sizeof(osb200_rayhit), "RayCastResult layout (GameMap.h:150-156)");
			EnsureInitialized();
			if (map) {
				std::lock_guard<std::mutex> guard(dirtyMutex);
				if (osb200_map_flush(ctx, solidMirror.data(), colorMirror.data()) < 0)
					RaiseAbi("osb200_map_flush");
			}
			if (osb200_castray2(ctx, reinterpret_cast<const float *>(v0), reinterpret_cast<const float *>(dir), maxSteps,
			                    reinterpret_cast<osb200_rayhit *>(out), n) != 0)
				RaiseAbi("osb200_castray2");
		}
	} // namespace draw
} // namespace spades
