/*
 * TEST INFRASTRUCTURE ONLY.  Drives the drop-in renderer (openspades_b200/host/ref_adapter/B200Renderer.cpp, compiled
 * here against the reference headers) and the unmodified draw::SWRenderer through ONE interface pointer type,
 * client::IRenderer*, on the same real client::GameMap: GameMap::Load of the VXL stream, SetGameMap (both register as
 * listeners), GameMap::Set edits (the real listener path fires in both), StartScene / RenderModel / AddLight /
 * EndScene / FrameDone / ReadBitmap / Flip -- the head-less pattern of Client/HitTestDebugger.cpp:38-57,147-305.
 * tests/test_gpu_adapter.py compares what the two ReadBitmap calls return.
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include <Client/GameMap.h>
#include <Client/IRenderer.h>
#include <Client/SceneDefinition.h>
#include <Core/Bitmap.h>
#include <Core/Debug.h>
#include <Core/Exception.h>
#include <Core/MemoryStream.h>
#include <Core/Settings.h>
#include <Core/VoxelModel.h>
#include <Draw/SWPort.h>
#include <Draw/SWRenderer.h>

#include "../../openspades_b200/host/ref_adapter/B200Renderer.h"

using namespace spades;

namespace {
	thread_local std::string g_err;

	class Port : public draw::SWPort { // HitTestDebugger::Port
		Handle<Bitmap> bmp[2];
		int cur = 0;

	public:
		int swaps = 0;
		Port(int w, int h) {
			bmp[0] = Handle<Bitmap>::New(w, h);
			bmp[1] = Handle<Bitmap>::New(w, h);
		}
		Bitmap &GetFramebuffer() override { return *bmp[cur]; }
		void Swap() override { // double-buffered like SDLSWPort: the pixel pointer changes every frame
			cur ^= 1;
			swaps++;
		}
	};

	struct Scene { // = osb200_scene / oref_scene
		float origin[3];
		float axis[3][3];
		float fovX, fovY, zNear, zFar;
	};

	client::SceneDefinition make_def(const Scene &s, int w, int h) {
		client::SceneDefinition def;
		def.viewportLeft = def.viewportTop = 0;
		def.viewportWidth = w;
		def.viewportHeight = h;
		def.fovX = s.fovX;
		def.fovY = s.fovY;
		def.viewOrigin = MakeVector3(s.origin[0], s.origin[1], s.origin[2]);
		for (int i = 0; i < 3; i++)
			def.viewAxis[i] = MakeVector3(s.axis[i][0], s.axis[i][1], s.axis[i][2]);
		def.zNear = s.zNear;
		def.zFar = s.zFar;
		def.skipWorld = false;
		return def;
	}

	Handle<VoxelModel> make_model(uint32_t seed) { // a small solid blob with a hole, seeded colours
		std::mt19937 rng(seed);
		const int w = 10, h = 12, d = 16;
		auto vm = Handle<VoxelModel>::New(w, h, d);
		for (int x = 0; x < w; x++)
			for (int y = 0; y < h; y++)
				for (int z = 0; z < d; z++) {
					const float dx = x - 4.5f, dy = y - 5.5f, dz = z - 7.5f;
					if (dx * dx / 25.f + dy * dy / 36.f + dz * dz / 64.f <= 1.f && !(x == 5 && y == 6))
						vm->SetSolid(x, y, z, (rng() & 0xffffffu) | (z == 3 ? 0u : 0x0u));
				}
		vm->SetOrigin(MakeVector3(-5.f, -6.f, -8.f));
		return vm;
	}
} // namespace

extern "C" {
__attribute__((visibility("default"))) const char *adapter_last_error(void) { return g_err.c_str(); }

/* Renders nScenes frames with both renderers; between frames editsPerFrame random GameMap::Set calls (dig / build) hit
 * the shared map.  withModels: every frame also submits 3 RenderModel calls and 2 AddLight calls.
 * outB200 / outSW: nScenes x h x w RGBA words as ReadBitmap returns them.  info: [0] listener calls seen by the B200
 * renderer's last flush (columns), [1] port swaps of the B200 port, [2] total edits applied.
 * numerics == 100: B200Renderer::NumericsOfThisBuild() against SWRenderer at its own default feature level (DetectFeatureLevel);
 * numerics < 0: only SWRenderer runs (usable without a GPU: the harness then serves to pin the oracle port on edited maps). */
__attribute__((visibility("default"))) int adapter_run(const void *vxl, size_t vxlLen, int w, int h, int numerics, int threads,
                                                       const void *scenesRaw, int nScenes, int editsPerFrame, int withModels,
                                                       uint32_t seed, uint32_t *outB200, uint32_t *outSW, int *info, int32_t *editsOut /* nullable: 6 ints per edit {frame, x, y, z, solid, colour} */) {
	// the game's main() does this before anything can throw: spades::Exception's constructor dereferences the
	// thread's backtrace object (Core/Exception.cpp:42, Core/Debug.cpp:64-73)
	reflection::Backtrace::StartBacktrace();
	try {
		{
			Settings::ItemHandle th("r_swNumThreads", nullptr);
			th = threads;
			Settings::ItemHandle un("r_swUndersampling", nullptr);
			un = 0;
		}
		const Scene *scenes = reinterpret_cast<const Scene *>(scenesRaw);
		MemoryStream ms(reinterpret_cast<const char *>(vxl), vxlLen);
		Handle<client::GameMap> map(client::GameMap::Load(&ms), false);

		auto portB = Handle<Port>::New(w, h), portS = Handle<Port>::New(w, h);
		// one interface type for both; the concrete classes appear only in these two lines
		Handle<client::IRenderer> rs[2];
		const int k0 = numerics < 0 ? 1 : 0;
		if (!k0)
			rs[0] = Handle<draw::B200Renderer>::New(portB.Cast<draw::SWPort>(),
			                                        numerics == 100 ? draw::B200Renderer::NumericsOfThisBuild() : numerics, 0)
			          .Cast<client::IRenderer>();
		rs[1] = Handle<draw::SWRenderer>::New(portS.Cast<draw::SWPort>(),
		                                      numerics == 100 ? draw::DetectFeatureLevel()
		                                                      : (numerics == 1 ? draw::SWFeatureLevel::SSE : draw::SWFeatureLevel::SSE2))
		          .Cast<client::IRenderer>();
		Handle<client::IModel> models[2];
		Handle<VoxelModel> vm = make_model(seed ^ 77u);
		for (int k = k0; k < 2; k++) {
			client::IRenderer &r = *rs[k];
			r.Init();
			r.SetFogColor(MakeVector3(0.8f, 1.f, 1.f));
			r.SetFogDistance(128.f);
			r.SetGameMap(*map);
			if (withModels)
				models[k] = r.CreateModel(*vm);
		}
		std::mt19937 rng(seed);
		int totalEdits = 0;
		for (int f = 0; f < nScenes; f++) {
			// edits through the real GameMap::Set: both renderers are listeners of the same map
			for (int e = 0; e < (f == 0 ? 0 : editsPerFrame); e++) {
				// in front of the camera (10..70 blocks along the view direction, +-15 sideways) so that the edits are
				// visible in the frame
				const float along = 10.f + (float)(rng() % 61), side = (float)(rng() % 31) - 15.f;
				const int x = (int)floorf(scenes[f].origin[0] + scenes[f].axis[2][0] * along + scenes[f].axis[0][0] * side) & 511;
				const int y = (int)floorf(scenes[f].origin[1] + scenes[f].axis[2][1] * along + scenes[f].axis[0][1] * side) & 511;
				const uint64_t col = map->GetSolidMapWrapped(x, y);
				int top = 0;
				while (top < 63 && !((col >> top) & 1ull))
					top++;
				int ez = -1, es = 0;
				uint32_t ec = 0;
				if ((rng() & 1u) && top < 61) {
					ez = top;
					map->Set(x, y, top, false, 0);
				} else if (top > 1) {
					ez = top - 1;
					es = 1;
					ec = 0x64000000u | (rng() & 0xffffffu);
					map->Set(x, y, top - 1, true, ec);
				}
				if (editsOut) {
					int32_t *e = editsOut + (size_t)totalEdits * 6;
					e[0] = f; e[1] = x; e[2] = y; e[3] = ez; e[4] = es; e[5] = (int32_t)ec;
				}
				totalEdits++;
			}
			const client::SceneDefinition def = make_def(scenes[f], w, h);
			for (int k = k0; k < 2; k++) {
				client::IRenderer &r = *rs[k];
				r.StartScene(def);
				if (withModels) {
					for (int i = 0; i < 3; i++) {
						client::ModelRenderParam p;
						const Vector3 at = def.viewOrigin + def.viewAxis[2] * (6.f + 5.f * i) + def.viewAxis[0] * (float)(i - 1) * 3.f;
						p.matrix = Matrix4::Translate(at) * Matrix4::Scale(0.25f + 0.1f * i);
						p.customColor = MakeVector3(0.2f * i, 0.5f, 1.f - 0.3f * i);
						r.RenderModel(*models[k], p);
					}
					for (int i = 0; i < 2; i++) {
						client::DynamicLightParam l;
						l.type = client::DynamicLightTypePoint;
						l.origin = def.viewOrigin + def.viewAxis[2] * (8.f + 6.f * i) + def.viewAxis[1] * 1.5f;
						l.radius = 9.f + 4.f * i;
						l.color = MakeVector3(1.f, 0.6f + 0.4f * i, 0.3f);
						r.AddLight(l);
					}
				}
				r.EndScene();
				r.FrameDone();
				Handle<Bitmap> bm = r.ReadBitmap();
				std::memcpy((k == 0 ? outB200 : outSW) + (size_t)f * w * h, bm->GetPixels(), sizeof(uint32_t) * w * h);
				r.Flip();
			}
		}
		if (info) {
			info[0] = k0 ? 0 : static_cast<draw::B200Renderer *>(rs[0].GetPointerOrNull())->LastFlushedColumns();
			info[1] = portB->swaps;
			info[2] = totalEdits;
		}
		// API misuse must raise like SWRenderer does (SWRenderer.cpp:520-547)
		int raised = 0;
		for (int k = k0; k < 2; k++) {
			try {
				rs[k]->EndScene();
			} catch (const spades::Exception &) {
				raised++;
			}
		}
		if (info)
			info[3] = raised;
		for (int k = k0; k < 2; k++) {
			models[k] = nullptr;
			rs[k]->SetGameMap({});
			rs[k]->Shutdown();
		}
		return 0;
	} catch (const std::exception &e) {
		g_err = e.what();
		return -1;
	}
}
}
