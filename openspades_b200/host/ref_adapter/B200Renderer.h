// The drop-in: a client::IRenderer + client::IGameMapListener for the reference tree, constructed from a
// Handle<draw::SWPort> exactly like draw::SWRenderer (Sources/Draw/SWRenderer.h:48-219, ctor :152), that sends the voxel
// hot path (terrain pass, voxel models, dynamic point lights, fog -- all of SWRenderer::EndScene up to the sprites,
// SWRenderer.cpp:709-745) to the B200 through the osb200 C ABI (include/osb200.h) instead of the CPU.
//
// This file is written against the reference's OWN headers (Client/IRenderer.h:98-174, Client/IGameMapListener.h:26-29,
// Draw/SWPort.h:28-43, Client/GameMap.h, Core/Bitmap.h ...) and is compiled only where a reference checkout exists
// (-I<openspades>/Sources); it is what a maintainer drops into Sources/Draw/ next to SWRenderer.cpp and selects in
// SDLRunner::CreateRenderer (Gui/SDLRunner.cpp:405-423).  oracle/Makefile target `adapter` compiles it in place with
// the reference translation units and tests/test_gpu_adapter.py drives it and the unmodified SWRenderer through the
// same client::IRenderer* calls (real GameMap::Load, real GameMap::Set -> listener path).
//
// Only the public interface of the reference types is used (GameMap::GetSolidMapWrapped / GetColor, Bitmap::GetPixels,
// VoxelModel::GetSolidBitsAt / GetColor ...): no friend declarations, no -fno-access-control.
#pragma once
#include <array>
#include <mutex>
#include <vector>

#include <Client/GameMap.h>
#include <Client/IGameMapListener.h>
#include <Client/IImage.h>
#include <Client/IModel.h>
#include <Client/IRenderer.h>
#include <Client/SceneDefinition.h>
#include <Core/Math.h>
#include <Draw/SWPort.h>

#include "../../../include/osb200.h"

namespace spades {
	class VoxelModel;
	namespace draw {

		/** client::IModel backed by a device copy of SWModel's render data (osb200_model). */
		class B200Model : public client::IModel {
			friend class B200Renderer;
			osb200_model *model;
			AABB3 bounds;

		protected:
			~B200Model();

		public:
			B200Model(osb200_model *m, const AABB3 &b) : model(m), bounds(b) {}
			AABB3 GetBoundingBox() override { return bounds; }
		};

		/** client::IImage placeholder: images feed sprites and the 2-D HUD (SWImageRenderer), which this renderer
		 * accepts and does not draw (out of the voxel hot path). */
		class B200Image : public client::IImage {
			float w, h;

		protected:
			~B200Image() {}

		public:
			B200Image(float w, float h) : w(w), h(h) {}
			void Update(Bitmap &, int, int) override {}
			float GetWidth() override { return w; }
			float GetHeight() override { return h; }
		};

		class B200Renderer : public client::IRenderer, public client::IGameMapListener {
			Handle<SWPort> port;
			Handle<client::GameMap> map;
			Bitmap *fb = nullptr; // owned by the port, re-fetched after every Swap (SWRenderer.cpp:1072-1075)
			osb200_ctx *ctx = nullptr;

			bool inited = false;
			bool duringSceneRendering = false;
			client::SceneDefinition sceneDef;
			float fogDistance = 128.f;
			Vector3 fogColor = MakeVector3(0.f, 0.f, 0.f);

			// host mirror of the map in GameMap's own array order, the source of osb200_map_flush's dirty columns
			std::vector<uint64_t> solidMirror; // [x][y]
			std::vector<uint32_t> colorMirror; // [x][y][z]
			std::mutex dirtyMutex;             // GameMapChanged may come from another thread (SWFlatMapRenderer.cpp:107-111)

			struct QueuedModel {
				Handle<B200Model> model; // keeps the model alive until EndScene, like SWRenderer::Model (SWRenderer.h:87-91)
				client::ModelRenderParam param;
			};
			std::vector<QueuedModel> models;
			std::vector<osb200_light> lights;

			void SetFramebuffer(Bitmap *);
			void EnsureInitialized();
			void EnsureValid();
			void EnsureSceneStarted();
			void EnsureSceneNotStarted();
			[[noreturn]] void RaiseAbi(const char *what);

		protected:
			~B200Renderer();

		public:
			/** numerics: OSB200_NUM_SSE2 reproduces SWRenderer at its default feature level on this host,
			 * OSB200_NUM_SSE / OSB200_NUM_IEEE the other two builds (SWFeatureLevel.h:25-121). */
			B200Renderer(Handle<SWPort> port, int numerics = OSB200_NUM_SSE2, int device = 0);

			/** the variant whose frames equal what SWRenderer draws in THIS binary on THIS host: the feature level
			 * SWRenderer's own constructor defaults to (DetectFeatureLevel, SWFeatureLevel.cpp:30-41) */
			static int NumericsOfThisBuild();

			void Init() override;
			void Shutdown() override;

			Handle<client::IImage> RegisterImage(const char *filename) override;
			Handle<client::IModel> RegisterModel(const char *filename) override;

			Handle<client::IImage> CreateImage(Bitmap &) override;
			Handle<client::IModel> CreateModel(VoxelModel &) override;

			void SetGameMap(stmp::optional<client::GameMap &>) override;
			void SetFogColor(Vector3 v) override;
			void SetFogDistance(float f) override { fogDistance = f; }

			void StartScene(const client::SceneDefinition &) override;

			void RenderModel(client::IModel &, const client::ModelRenderParam &) override;
			void AddLight(const client::DynamicLightParam &light) override;

			void AddDebugLine(Vector3 a, Vector3 b, Vector4 color) override;
			void AddSprite(client::IImage &, Vector3 center, float radius, float rotation) override;
			void AddLongSprite(client::IImage &, Vector3 p1, Vector3 p2, float radius) override;

			void EndScene() override;

			void MultiplyScreenColor(Vector3) override;
			void SetColor(Vector4) override;
			void SetColorAlphaPremultiplied(Vector4) override;

			void DrawImage(stmp::optional<client::IImage &>, const Vector2 &outTopLeft) override;
			void DrawImage(stmp::optional<client::IImage &>, const AABB2 &outRect) override;
			void DrawImage(stmp::optional<client::IImage &>, const Vector2 &outTopLeft, const AABB2 &inRect) override;
			void DrawImage(stmp::optional<client::IImage &>, const AABB2 &outRect, const AABB2 &inRect) override;
			void DrawImage(stmp::optional<client::IImage &>, const Vector2 &outTopLeft, const Vector2 &outTopRight,
			               const Vector2 &outBottomLeft, const AABB2 &inRect) override;

			void DrawFlatGameMap(const AABB2 &outRect, const AABB2 &inRect) override;

			void FrameDone() override;
			void Flip() override;
			Handle<Bitmap> ReadBitmap() override;

			float ScreenWidth() override;
			float ScreenHeight() override;

			void GameMapChanged(int x, int y, int z, client::GameMap *) override;

			/** SWRenderer::depthBuffer of the last frame (view-space Z), copied from the device. */
			std::vector<float> ReadDepthBuffer();
			/** columns re-uploaded by the last EndScene (dirty columns only) */
			int LastFlushedColumns() const { return lastFlushed; }
			/** GameMap::CastRay2 for a batch of rays on the resident map (Client/GameMap.cpp:361-507) */
			void CastRay2Batch(const Vector3 *v0, const Vector3 *dir, int maxSteps, client::GameMap::RayCastResult *out,
			                   size_t n);

		private:
			int lastFlushed = 0;
		};
	} // namespace draw
} // namespace spades
