// Host-side renderer with draw::SWRenderer's public surface for the voxel hot path
// (Sources/Draw/SWRenderer.h:48-219 implementing client::IRenderer, Sources/Client/IRenderer.h:98-174, and
// client::IGameMapListener, Sources/Client/IGameMapListener.h:26-29).  Same method names, argument meaning,
// call-order rules and error messages; every piece of per-pixel / per-voxel work goes through the osb200 C ABI
// (include/osb200.h) to the CUDA kernels.  There is no CPU rendering path in this class.
//
// What is mirrored:  Init / Shutdown, SetGameMap (+ listener registration), SetFogDistance / SetFogColor,
//   StartScene / EndScene (terrain pass, voxel models, dynamic point lights and fog of EndScene, SWRenderer.cpp:709-745),
//   CreateModel / RenderModel (SWModel + SWModelRenderer), AddLight (SWRenderer.cpp:578-649), FrameDone, Flip, ReadBitmap,
//   ScreenWidth / ScreenHeight, GameMapChanged -> dirty-column bitset -> re-upload at the next EndScene
//   (SWFlatMapRenderer::SetNeedsUpdate / Update, SWFlatMapRenderer.cpp:52-111).
// What is accepted and NOT drawn (out of scope): AddSprite, AddLongSprite, AddDebugLine and the
//   2-D DrawImage family; they only enforce the same call-order rules.
// Extension beyond SWRenderer (the CPU halves of GLRenderer's map shadow + radiosity):
//   UpdateRadiosity(), and of the ray-traced ambient occlusion volume (GLAmbientShadowRenderer): UpdateAmbient().
//
// The reference's own types (Vector3, Handle<T>, Bitmap, SWPort) are not linkable here, so this header names
// small layout-compatible stand-ins; INTEGRATION.md shows the 1:1 adapter a maintainer adds inside Sources/Draw.
#pragma once
#include <cstdint>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <vector>

#include "../../include/osb200.h"
#include "VoxelMap.h"

namespace osb200 {

	struct Vector3 {
		float x, y, z;
	};

	// client::SceneDefinition (Sources/Client/SceneDefinition.h:26-46); only the geometric fields reach the kernels
	struct SceneDefinition {
		int viewportLeft = 0, viewportTop = 0, viewportWidth = 0, viewportHeight = 0;
		float fovX = 0.f, fovY = 0.f;
		Vector3 viewOrigin{0, 0, 0};
		Vector3 viewAxis[3]{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}; // right, up, front
		float zNear = 0.1f, zFar = 200.f;
		bool skipWorld = false;
		float depthOfFieldFocalLength = 0.f, depthOfFieldNearBlurStrength = 0.f, depthOfFieldFarBlurStrength = 0.f;
		unsigned int time = 0;
		bool denyCameraBlur = false;
		float blurVignette = 0.f, globalBlur = 0.f, saturation = 1.f, radialBlur = 0.f;
	};

	// spades::Bitmap as the SW port hands it out: BGRX words, top-down rows
	class Framebuffer {
	public:
		Framebuffer(int w, int h) : w_(w), h_(h), px_((size_t)w * h, 0u) {}
		int GetWidth() const { return w_; }
		int GetHeight() const { return h_; }
		uint32_t *GetPixels() { return px_.data(); }
		const uint32_t *GetPixels() const { return px_.data(); }

	private:
		int w_, h_;
		std::vector<uint32_t> px_;
	};

	// draw::SWPort (Sources/Draw/SWPort.h:28-43): the head-less framebuffer provider
	class FramebufferPort {
	public:
		virtual ~FramebufferPort() {}
		virtual Framebuffer &GetFramebuffer() = 0;
		virtual void Swap() = 0;
	};

	// in-memory port (the pattern of Client/HitTestDebugger.cpp:38-50)
	class MemoryPort : public FramebufferPort {
	public:
		MemoryPort(int w, int h) : fb_(w, h) {}
		Framebuffer &GetFramebuffer() override { return fb_; }
		void Swap() override { swaps_++; }
		int SwapCount() const { return swaps_; }

	private:
		Framebuffer fb_;
		int swaps_ = 0;
	};

	// what SPRaise throws in the reference (Core/Exception.h:29-46)
	class RendererError : public std::runtime_error {
	public:
		explicit RendererError(const std::string &m) : std::runtime_error(m) {}
	};

	struct B200RendererOptions {
		int device = 0;
		int numerics = OSB200_NUM_SSE2; // the reference's default feature level (SWFeatureLevel.cpp:29-37)
		int swThreads = 4;              // r_swNumThreads (SWRenderer.cpp:41)
		int undersampling = 1;          // r_swUndersampling (SWMapRenderer.cpp:37)
		int radiosityMode = 1;          // r_radiosity
	};

	// spades::VoxelModel's data as SWModel reads it (Core/VoxelModel.h:68-106): solid[x + y*w] bit z, colors[(x + y*w)*d + z]
	struct VoxelModelData {
		int width = 0, height = 0, depth = 0;
		Vector3 origin{0, 0, 0};
		std::vector<uint64_t> solidBits;
		std::vector<uint32_t> colors;
	};
	// client::ModelRenderParam (IRenderer.h:39-63): the two fields the software renderer reads
	struct ModelRenderParam {
		float matrix[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}; // Matrix4::m, column-major
		Vector3 customColor{0, 0, 0};
	};
	enum DynamicLightType { DynamicLightTypePoint, DynamicLightTypeSpotlight, DynamicLightTypeLinear };
	// client::DynamicLightParam (IRenderer.h:71-96); only point lights are drawn (SWRenderer.cpp:581-585)
	struct DynamicLightParam {
		DynamicLightType type = DynamicLightTypePoint;
		Vector3 origin{0, 0, 0};
		float radius = 1.f;
		Vector3 color{1, 1, 1};
	};
	class B200Renderer;
	// client::IModel as SWRenderer::CreateModel returns it (draw::SWModel): owns the device copy of the render data
	class Model {
	public:
		~Model();
		Model(const Model &) = delete;
		Model &operator=(const Model &) = delete;

	private:
		friend class B200Renderer;
		Model(osb200_model *m, const B200Renderer *owner) : m_(m, osb200_model_destroy), owner_(owner) {}
		// shared with the renderer's queue: a model destroyed between RenderModel and EndScene stays alive until the
		// scene has been drawn, like the Handle<SWModel> SWRenderer keeps (SWRenderer.h:87-91)
		std::shared_ptr<osb200_model> m_;
		const B200Renderer *owner_;
	};

	class B200Renderer : public IVoxelMapListener {
	public:
		explicit B200Renderer(std::shared_ptr<FramebufferPort> port, const B200RendererOptions &opts = B200RendererOptions());
		~B200Renderer() override;

		void Init();
		void Shutdown();

		void SetGameMap(VoxelMap *map); // nullptr detaches (stmp::optional<GameMap&> in the reference)
		// GameMap::Load + SetGameMap in one step: the raw VXL stream is decoded on the device (osb200_map_load_vxl), the
		// returned VoxelMap receives the decoded arrays and becomes the current map without a second upload.
		// Throws RendererError on a truncated stream (the reference's loader throws as well).
		std::unique_ptr<VoxelMap> LoadGameMap(const uint8_t *data, size_t len);
		// GameMap::Save of the current map, encoded on the device (pending edits are flushed first)
		std::vector<uint8_t> SaveGameMap();
		void SetFogDistance(float v) { fogDistance_ = v; }
		void SetFogColor(Vector3 v) { fogColor_ = v; }

		void StartScene(const SceneDefinition &def);
		// SWRenderer::CreateModel (SWRenderer.cpp:142-147): builds SWModel's render data and uploads it
		std::unique_ptr<Model> CreateModel(const VoxelModelData &vm);
		// queued until EndScene, drawn in submission order between terrain and fog
		void AddLight(const DynamicLightParam &param);
		void RenderModel(Model &model, const ModelRenderParam &param);
		// accepted, order-checked, not drawn (see header comment)
		void AddDebugLine(Vector3 a, Vector3 b, const float color[4]);
		void AddSprite(const void *image, Vector3 center, float radius, float rotation);
		void AddLongSprite(const void *image, Vector3 p1, Vector3 p2, float radius);
		void EndScene();

		void FrameDone();
		void Flip();
		// RGBA, bottom-up copy of the framebuffer with alpha forced to 0xff (SWRenderer.cpp:1078-1098)
		std::unique_ptr<Framebuffer> ReadBitmap();

		float ScreenWidth() const;
		float ScreenHeight() const;

		// client::IGameMapListener
		void GameMapChanged(int x, int y, int z, VoxelMap *map) override;

		// view-space Z of the last frame (SWRenderer::depthBuffer), copied from the device on demand
		std::vector<float> ReadDepthBuffer();

		// GLRenderer::EndScene's "Upload Dynamic Data" leg for the map (GLRenderer.cpp:745-759):
		// GLMapShadowRenderer::Update + GLRadiosityRenderer::Update with a chunk budget (8 in the reference,
		// < 0 = drain); prioritises chunks near the last scene's eye.  Returns the number of chunks updated.
		int UpdateRadiosity(int maxChunks = 8);
		// Same leg for the ray-traced ambient-occlusion volume (GLAmbientShadowRenderer::Update ->
		// UpdateDirtyChunks -> UpdateChunk, GLAmbientShadowRenderer.cpp:239-507; r_radiosity's sibling cvar
		// r_ambientOcclusion... the reference runs it on a worker thread, 8 chunks per dispatch).
		int UpdateAmbient(int maxChunks = 8);
		// the 16 ray directions of the ambient evaluation (GLAmbientShadowRenderer ctor :53-60 draws them at random)
		void SetAmbientRays(const float rays[48]);

		osb200_ctx *Context() { return ctx_; }
		int LastFlushedColumns() const { return lastFlushed_; }

	private:
		void EnsureInitialized() const;
		void EnsureValid() const;
		void EnsureSceneStarted() const;
		void EnsureSceneNotStarted() const;
		void SetFramebuffer(Framebuffer *fb);
		[[noreturn]] void RaiseAbi(const char *what) const;

		std::shared_ptr<FramebufferPort> port_;
		B200RendererOptions opts_;
		osb200_ctx *ctx_ = nullptr;
		VoxelMap *map_ = nullptr;
		Framebuffer *fb_ = nullptr;
		bool inited_ = false;
		bool duringSceneRendering_ = false;
		bool haveScene_ = false;
		SceneDefinition sceneDef_;
		float fogDistance_ = 128.f;
		Vector3 fogColor_{0.f, 0.f, 0.f};
		int lastFlushed_ = 0;
		std::vector<osb200_model_instance> models_; // RenderModel calls of the open scene
		std::vector<std::shared_ptr<osb200_model>> modelRefs_; // ... and what keeps their models alive until EndScene
		std::vector<osb200_light> lights_;          // AddLight calls of the open scene (point lights)
		std::mutex dirtyMutex_; // GameMapChanged may arrive from another thread (SWFlatMapRenderer.cpp:107-111)
	};
} // namespace osb200
