// See B200Renderer.h.  Reference behaviour followed, method by method: Sources/Draw/SWRenderer.cpp.
#include "B200Renderer.h"

#include <cstring>

namespace osb200 {

	B200Renderer::B200Renderer(std::shared_ptr<FramebufferPort> port, const B200RendererOptions &opts)
	    : port_(std::move(port)), opts_(opts) {
		// SWRenderer::SWRenderer (SWRenderer.cpp:45-87): needs a port, takes its framebuffer, allocates depth
		if (!port_)
			throw RendererError("Renderer is not valid");
		SetFramebuffer(&port_->GetFramebuffer());
		osb200_opts o;
		std::memset(&o, 0, sizeof(o));
		o.device = opts_.device;
		o.numerics = opts_.numerics;
		o.sw_threads = opts_.swThreads;
		o.undersampling = opts_.undersampling;
		o.radiosity_mode = opts_.radiosityMode;
		if (osb200_create(&o, &ctx_) != 0)
			throw RendererError(std::string("B200Renderer: ") + osb200_last_error());
	}

	B200Renderer::~B200Renderer() {
		Shutdown();
		if (ctx_)
			osb200_destroy(ctx_);
		ctx_ = nullptr;
	}

	void B200Renderer::RaiseAbi(const char *what) const {
		throw RendererError(std::string(what) + ": " + osb200_last_error());
	}

	void B200Renderer::SetFramebuffer(Framebuffer *fb) {
		// SWRenderer.cpp:94-108
		if (fb == nullptr)
			throw RendererError("Framebuffer is null.");
		if (fb_ && (fb->GetWidth() != fb_->GetWidth() || fb->GetHeight() != fb_->GetHeight()))
			throw RendererError("Framebuffer size changed."); // SPAssert in the reference
		fb_ = fb;
		if ((fb->GetWidth() & 7) || (fb->GetHeight() & 7))
			throw RendererError("Framebuffer size is not multiple of 8.");
	}

	void B200Renderer::Init() { inited_ = true; } // SWRenderer.cpp:110-124 (model renderer: out of scope)

	void B200Renderer::Shutdown() {
		// SWRenderer.cpp:126-140
		if (map_)
			SetGameMap(nullptr);
		port_.reset();
		fb_ = nullptr;
		inited_ = false;
	}

	void B200Renderer::EnsureSceneStarted() const {
		if (!duringSceneRendering_)
			throw RendererError("Illegal call outside of StartScene ... EndScene");
	}
	void B200Renderer::EnsureSceneNotStarted() const {
		if (duringSceneRendering_)
			throw RendererError("Illegal call between StartScene ... EndScene");
	}
	void B200Renderer::EnsureInitialized() const {
		if (!inited_)
			throw RendererError("Renderer is not initialized");
		EnsureValid();
	}
	void B200Renderer::EnsureValid() const {
		if (!port_)
			throw RendererError("Renderer is not valid");
	}

	void B200Renderer::SetGameMap(VoxelMap *map) {
		// SWRenderer.cpp:172-191
		if (map)
			EnsureInitialized();
		if (map == map_)
			return;
		if (map_)
			map_->RemoveListener(this);
		map_ = map;
		if (!map_)
			osb200_map_clear(ctx_);
		if (map_) {
			if (map_->Width() != OSB200_MAP_W || map_->Height() != OSB200_MAP_H || map_->Depth() != OSB200_MAP_D)
				throw RendererError("B200Renderer: map must be 512x512x64"); // SWMapRenderer.cpp:224-225
			map_->AddListener(this);
			// SWMapRenderer ctor (RLE build, SWMapRenderer.cpp:107-138) == pack + upload of the whole map
			if (osb200_map_upload(ctx_, map_->SolidData(), map_->ColorData()) != 0)
				RaiseAbi("osb200_map_upload");
		}
	}

	std::unique_ptr<VoxelMap> B200Renderer::LoadGameMap(const uint8_t *data, size_t len) {
		EnsureInitialized();
		EnsureSceneNotStarted();
		std::unique_ptr<VoxelMap> map(new VoxelMap());
		if (map_)
			map_->RemoveListener(this);
		map_ = nullptr;
		if (osb200_map_load_vxl(ctx_, data, len, map->MutableSolidData(), map->MutableColorData()) != 0) {
			osb200_map_clear(ctx_);
			RaiseAbi("osb200_map_load_vxl");
		}
		map_ = map.get();
		map_->AddListener(this);
		return map;
	}

	std::vector<uint8_t> B200Renderer::SaveGameMap() {
		EnsureInitialized();
		EnsureSceneNotStarted();
		if (!map_)
			throw RendererError("B200Renderer: no map set");
		{
			std::lock_guard<std::mutex> guard(dirtyMutex_);
			lastFlushed_ = osb200_map_flush(ctx_, map_->SolidData(), map_->ColorData());
		}
		if (lastFlushed_ < 0)
			RaiseAbi("osb200_map_flush");
		const long long n = osb200_map_save_vxl(ctx_, nullptr, 0);
		if (n < 0)
			RaiseAbi("osb200_map_save_vxl");
		std::vector<uint8_t> out((size_t)n);
		if (osb200_map_save_vxl(ctx_, out.data(), out.size()) != n)
			RaiseAbi("osb200_map_save_vxl");
		return out;
	}

	void B200Renderer::GameMapChanged(int x, int y, int z, VoxelMap *map) {
		// SWRenderer.cpp:1148-1154 -> SWFlatMapRenderer::SetNeedsUpdate (mutex-protected bitset)
		if (map != map_)
			return;
		std::lock_guard<std::mutex> guard(dirtyMutex_);
		osb200_map_mark_dirty(ctx_, x, y, z);
	}

	void B200Renderer::StartScene(const SceneDefinition &def) {
		// SWRenderer.cpp:549-563
		EnsureInitialized();
		EnsureSceneNotStarted();
		sceneDef_ = def;
		haveScene_ = true;
		duringSceneRendering_ = true;
	}

	Model::~Model() = default; // the device copy goes with the last reference (this one, or the open scene's queue)

	std::unique_ptr<Model> B200Renderer::CreateModel(const VoxelModelData &vm) {
		EnsureInitialized();
		if (vm.width <= 0 || vm.height <= 0 || vm.depth <= 0 ||
		    vm.solidBits.size() != (size_t)vm.width * vm.height ||
		    vm.colors.size() != (size_t)vm.width * vm.height * vm.depth)
			throw RendererError("invalid voxel model");
		const float origin[3] = {vm.origin.x, vm.origin.y, vm.origin.z};
		osb200_model *m = nullptr;
		if (osb200_model_create(ctx_, vm.width, vm.height, vm.depth, origin, vm.solidBits.data(), vm.colors.data(), &m) != 0)
			RaiseAbi("osb200_model_create");
		return std::unique_ptr<Model>(new Model(m, this));
	}

	void B200Renderer::AddLight(const DynamicLightParam &param) {
		// SWRenderer.cpp:578-585; the culls and the screen rectangle (:587-649) are evaluated inside osb200_render_scene
		EnsureInitialized();
		EnsureSceneStarted();
		if (param.type != DynamicLightTypePoint)
			return;
		osb200_light l;
		l.origin[0] = param.origin.x;
		l.origin[1] = param.origin.y;
		l.origin[2] = param.origin.z;
		l.radius = param.radius;
		l.color[0] = param.color.x;
		l.color[1] = param.color.y;
		l.color[2] = param.color.z;
		lights_.push_back(l);
	}
	void B200Renderer::RenderModel(Model &model, const ModelRenderParam &param) {
		// SWRenderer.cpp:565-576
		EnsureInitialized();
		EnsureSceneStarted();
		if (model.owner_ != this)
			throw RendererError("model"); // SPInvalidArgument("model"): the dynamic_cast<SWModel&> of a foreign model
		osb200_model_instance inst;
		inst.model = model.m_.get();
		modelRefs_.push_back(model.m_);
		for (int i = 0; i < 16; i++)
			inst.matrix[i] = param.matrix[i];
		inst.customColor[0] = param.customColor.x;
		inst.customColor[1] = param.customColor.y;
		inst.customColor[2] = param.customColor.z;
		models_.push_back(inst);
	}
	void B200Renderer::AddDebugLine(Vector3, Vector3, const float *) {
		EnsureInitialized();
		EnsureSceneStarted();
	}
	void B200Renderer::AddSprite(const void *, Vector3, float, float) {
		EnsureInitialized();
		EnsureSceneStarted();
	}
	void B200Renderer::AddLongSprite(const void *, Vector3, Vector3, float) {
		EnsureInitialized();
		EnsureSceneStarted();
	}

	void B200Renderer::EndScene() {
		// SWRenderer.cpp:709-745: clear to 0x7f7f7f, flat-map Update (drains dirty columns -> UpdateRle x5),
		// terrain, models, lights, fog.  The clear is folded into the kernels: the terrain
		// pass writes every pixel, or 0x7f7f7f when the eye is inside a solid voxel (SWMapRenderer.cpp:1181-1184).
		EnsureInitialized();
		EnsureSceneStarted();
		duringSceneRendering_ = false;
		const int w = fb_->GetWidth(), h = fb_->GetHeight();
		// Without a map the reference still clears and fogs over the stale depth buffer (mapRenderer == null,
		// :719-744); osb200_render does exactly that when no map is resident.  SceneDefinition::skipWorld is
		// not consulted by SWRenderer either.
		if (map_) {
			{
				std::lock_guard<std::mutex> guard(dirtyMutex_);
				lastFlushed_ = osb200_map_flush(ctx_, map_->SolidData(), map_->ColorData());
			}
			if (lastFlushed_ < 0)
				RaiseAbi("osb200_map_flush");
		}
		osb200_scene sc;
		sc.origin[0] = sceneDef_.viewOrigin.x;
		sc.origin[1] = sceneDef_.viewOrigin.y;
		sc.origin[2] = sceneDef_.viewOrigin.z;
		for (int i = 0; i < 3; i++) {
			sc.axis[i][0] = sceneDef_.viewAxis[i].x;
			sc.axis[i][1] = sceneDef_.viewAxis[i].y;
			sc.axis[i][2] = sceneDef_.viewAxis[i].z;
		}
		sc.fovX = sceneDef_.fovX;
		sc.fovY = sceneDef_.fovY;
		sc.zNear = sceneDef_.zNear;
		sc.zFar = sceneDef_.zFar;
		const float fog[3] = {fogColor_.x, fogColor_.y, fogColor_.z};
		// models.clear() / lights.clear() of the reference (:731,737) happen whether or not the draw succeeds
		std::vector<osb200_model_instance> models;
		std::vector<osb200_light> lights;
		std::vector<std::shared_ptr<osb200_model>> refs;
		models.swap(models_);
		refs.swap(modelRefs_);
		lights.swap(lights_);
		if (osb200_render_scene(ctx_, &sc, w, h, fog, fogDistance_, models.data(), (int)models.size(), lights.data(),
		                        (int)lights.size(), fb_->GetPixels(), nullptr) != 0)
			RaiseAbi("osb200_render_scene");
	}

	void B200Renderer::FrameDone() {
		EnsureValid();
		EnsureSceneNotStarted();
	}

	void B200Renderer::Flip() {
		// SWRenderer.cpp:1045-1076: the pixel pointer may change every frame
		EnsureValid();
		EnsureSceneNotStarted();
		port_->Swap();
		SetFramebuffer(&port_->GetFramebuffer());
	}

	std::unique_ptr<Framebuffer> B200Renderer::ReadBitmap() {
		EnsureValid();
		EnsureSceneNotStarted();
		const int w = fb_->GetWidth(), h = fb_->GetHeight();
		std::unique_ptr<Framebuffer> bm(new Framebuffer(w, h));
		const uint32_t *in = fb_->GetPixels();
		uint32_t *out = bm->GetPixels();
		for (int y = 0; y < h; y++) {
			const uint32_t *src = in + (size_t)y * w;
			uint32_t *dst = out + (size_t)(h - 1 - y) * w;
			for (int x = 0; x < w; x++) {
				const uint32_t c = src[x];
				dst[x] = 0xff000000u | (c & 0xff00u) | ((c & 0xffu) << 16) | ((c & 0xff0000u) >> 16);
			}
		}
		return bm;
	}

	float B200Renderer::ScreenWidth() const {
		EnsureValid();
		return static_cast<float>(fb_->GetWidth());
	}
	float B200Renderer::ScreenHeight() const {
		EnsureValid();
		return static_cast<float>(fb_->GetHeight());
	}

	std::vector<float> B200Renderer::ReadDepthBuffer() {
		EnsureValid();
		const int w = fb_->GetWidth(), h = fb_->GetHeight();
		std::vector<float> d((size_t)w * h, 0.f);
		if (osb200_read_depth(ctx_, d.data(), d.size()) != 0)
			RaiseAbi("osb200_read_depth");
		return d;
	}

	int B200Renderer::UpdateRadiosity(int maxChunks) {
		EnsureInitialized();
		EnsureSceneNotStarted();
		if (!map_)
			return 0;
		{
			std::lock_guard<std::mutex> guard(dirtyMutex_);
			lastFlushed_ = osb200_map_flush(ctx_, map_->SolidData(), map_->ColorData());
		}
		if (lastFlushed_ < 0)
			RaiseAbi("osb200_map_flush");
		const float eye[3] = {sceneDef_.viewOrigin.x, sceneDef_.viewOrigin.y, sceneDef_.viewOrigin.z};
		const int n = osb200_radiosity_update(ctx_, maxChunks, haveScene_ ? eye : nullptr);
		if (n < 0)
			RaiseAbi("osb200_radiosity_update");
		return n;
	}
	int B200Renderer::UpdateAmbient(int maxChunks) {
		EnsureInitialized();
		EnsureSceneNotStarted();
		if (!map_)
			return 0;
		{
			std::lock_guard<std::mutex> guard(dirtyMutex_);
			lastFlushed_ = osb200_map_flush(ctx_, map_->SolidData(), map_->ColorData());
		}
		if (lastFlushed_ < 0)
			RaiseAbi("osb200_map_flush");
		const float eye[3] = {sceneDef_.viewOrigin.x, sceneDef_.viewOrigin.y, sceneDef_.viewOrigin.z};
		const int n = osb200_ambient_update(ctx_, maxChunks, haveScene_ ? eye : nullptr);
		if (n < 0)
			RaiseAbi("osb200_ambient_update");
		return n;
	}

	void B200Renderer::SetAmbientRays(const float rays[48]) {
		EnsureInitialized();
		if (osb200_ambient_set_rays(ctx_, rays) != 0)
			RaiseAbi("osb200_ambient_set_rays");
	}
} // namespace osb200
