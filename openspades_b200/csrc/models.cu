// Voxel models and dynamic lights of the software renderer's EndScene, between the terrain pass and the fog
// (SWRenderer::EndScene, Sources/Draw/SWRenderer.cpp:709-745):
//   SWModel::SWModel                  Sources/Draw/SWModel.cpp:25-100         (build_render_data, host, once per model)
//   SWModelRenderer::RenderInner      Sources/Draw/SWModelRenderer.cpp:46-277 (k_model_setup, k_model_splat; winners resolved in
//                                                                              terrain.cu k_scene_finish)
//   SWRenderer::StartScene matrices   Sources/Client/SceneDefinition.cpp:26-72, Sources/Core/Math.cpp:108-181 (host)
//   SWRenderer::AddLight              Sources/Draw/SWRenderer.cpp:578-649     (host: culls + screen rectangle)
//   SWRenderer::ApplyDynamicLight     Sources/Draw/SWRenderer.cpp:229-339     (k_light_prep here, the per-pixel part in k_scene_finish)
//
// Design: the reference splats every surface voxel as a screen-aligned square with a strict depth test, models in
// submission order, voxels in x / y / z order -- i.e. per pixel the winner is the smallest depth and, among equal
// depths, the earliest splat; the terrain wins ties against every splat.  That is one 64-bit atomicMin per covered
// pixel on the key (order-preserving depth bits << 32 | sequence number), with the terrain entered as sequence 0,
// followed by a resolve pass that fetches the winner's colour.  One warp per (instance, model column); the
// reference's sequential float recurrences (v1 += tAxis1 per x, v2 += tAxis2 per y) are replayed per thread.
// Lights modify a pixel in submission order, so one thread per pixel walks the light list; the per-row / per-column
// float recurrences of the reference (and its per-thread row bands) come from small tables built by one thread per light.
#include "ctx.h"
#include "numerics.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

struct osb200_model {
	int w, h, d;
	float origin[3]; // VoxelModel::GetOrigin
	float center[3]; // SWModel::center = dims * 0.5 (SWModel.cpp:28-32)
	float radius;    // |center|
	std::vector<uint32_t> renderData, renderDataAddr; // host copies (test hook)
	uint32_t *d_data = nullptr, *d_addr = nullptr;
	uint32_t voxels = 0; // splat candidates = (renderData.size() - columns) / 2
	osb200_ctx *owner = nullptr; // identity check only (osb200_render_scene); never dereferenced through a model
	int device = 0;              // the owner's device, kept here so that a model may outlive its context
};

namespace osb {

	// SWModel ctor (SWModel.cpp:34-99): per column the surface voxels as (z << 24 | BGR-swizzled colour, normal index)
	// pairs, terminated by 0xffffffff; columns are appended y-major, so column (x, y) starts at addr[x + y * w].
	static void build_render_data(osb200_model &M, const uint64_t *solid, const uint32_t *colors) {
		const int w = M.w, h = M.h, d = M.d;
		auto bits = [&](int x, int y) -> uint64_t { return solid[x + y * w]; };
		auto isSolid = [&](int x, int y, int z) -> bool { // VoxelModel::IsSolid (VoxelModel.h:126-132)
			if (x < 0 || y < 0 || x >= w || y >= h || z < 0 || z >= d)
				return false;
			return (bits(x, y) >> z) & 1;
		};
		for (int y = 0; y < h; y++)
			for (int x = 0; x < w; x++) {
				M.renderDataAddr.push_back((uint32_t)M.renderData.size());
				const uint64_t map = bits(x, y);
				uint64_t around = x > 0 ? bits(x - 1, y) : 0;
				around &= x < w - 1 ? bits(x + 1, y) : 0;
				around &= y > 0 ? bits(x, y - 1) : 0;
				around &= y < h - 1 ? bits(x, y + 1) : 0;
				for (int z = 0; z < d; z++) {
					if (!(map & (1ULL << z)))
						continue;
					// surface test: top / bottom layer, a hole above or below, or an open side
					if (!(z == 0 || z == d - 1 || ((map >> (z - 1)) & 7ULL) != 7ULL || (around & (1ULL << z)) == 0))
						continue;
					const uint32_t col = colors[(size_t)(x + y * w) * d + z];
					M.renderData.push_back(((col & 0xff00) | ((col & 0xff) << 16) | ((col & 0xff0000) >> 16)) | ((uint32_t)z << 24));
					uint32_t normal;
					if ((col >> 24) == 1) { // MaterialType::Emissive
						normal = 27;
					} else {
						int n[3] = {0, 0, 0};
						for (int cx = -1; cx <= 1; cx++)
							for (int cy = -1; cy <= 1; cy++)
								for (int cz = -1; cz <= 1; cz++) {
									const int sgn = isSolid(x + cx, y + cy, z + cz) ? -1 : 1;
									n[0] += sgn * cx;
									n[1] += sgn * cy;
									n[2] += sgn * cz;
								}
						for (int k = 0; k < 3; k++)
							n[k] = std::max(std::min(n[k], 1), -1) + 1;
						normal = (uint32_t)(n[0] + n[1] * 3 + n[2] * 9);
					}
					M.renderData.push_back(normal);
					M.voxels++;
				}
				M.renderData.push_back(0xffffffffU);
			}
	}

	struct ModelDev { // device-side description of one uploaded model
		const uint32_t *data, *addr;
		int w, h;
		float origin[3], center[3], radius;
	};

	struct InstanceIn { // host -> device, per RenderModel call
		ModelDev model;
		float matrix[16];
		uint32_t customColor; // ToFixed8 b | g << 8 | r << 16 (SWModelRenderer.cpp:180-182)
		uint32_t seqBase;     // sequence number of this instance's first voxel (1-based; 0 is the terrain)
		uint32_t colBase;     // first (instance, column) work item
	};

	struct InstanceSetup { // k_model_setup -> k_model_splat
		float tOrigin[4], tAxis1[4], tAxis2[4], tAxis3[4];
		float pointDiameter;
		int visible;
		uint8_t brights[28];
	};

	struct SceneMats { // per scene, computed on the host with the reference's float operations
		float viewproj[16];
		float frustum[6][4]; // n.xyz, w
		float zNear;
		int fw, fh;
	};

	struct ModelState {
		InstanceIn *d_inst = nullptr, *h_inst = nullptr;
		InstanceSetup *d_setup = nullptr;
		size_t instCap = 0, dInstCap = 0, dSetupCap = 0, dLightCap = 0;
		SceneMats *d_mats = nullptr, *h_mats = nullptr;
		unsigned long long *d_keys = nullptr;
		size_t keyCap = 0;
		uint32_t *d_splatColor = nullptr;
		size_t splatCap = 0;
		LightDev *d_lights = nullptr, *h_lights = nullptr;
		size_t lightCap = 0;
		float *d_lightTabs = nullptr;
		size_t lightTabCap = 0;
		cudaEvent_t stageFree = nullptr;
		cudaStream_t aux = nullptr;   // per-instance set-up and light tables run beside the terrain kernels
		cudaEvent_t auxDone = nullptr;
	};

	// ---- small float helpers in the reference's operation order (Core/Math.h, Core/Math.cpp:162-181) ----
	__device__ __forceinline__ float dot3(const float a[3], const float b[3]) {
		return __fadd_rn(__fadd_rn(__fmul_rn(a[0], b[0]), __fmul_rn(a[1], b[1])), __fmul_rn(a[2], b[2]));
	}
	__device__ __forceinline__ float len2_3(const float a[3]) { return dot3(a, a); }
	__device__ __forceinline__ void madd3(float o[3], const float a[3], float s) { // o += a * s
#pragma unroll
		for (int k = 0; k < 3; k++)
			o[k] = __fadd_rn(o[k], __fmul_rn(a[k], s));
	}
	__device__ __forceinline__ void mat_vec4(const float *m, const float v[3], float w, float out[4]) {
#pragma unroll
		for (int r = 0; r < 4; r++) {
			float acc = __fmul_rn(m[r], v[0]);
			acc = __fadd_rn(acc, __fmul_rn(m[4 + r], v[1]));
			acc = __fadd_rn(acc, __fmul_rn(m[8 + r], v[2]));
			acc = __fadd_rn(acc, __fmul_rn(m[12 + r], w));
			out[r] = acc;
		}
	}
	__device__ __forceinline__ float fmax_std(float a, float b) { return a < b ? b : a; } // std::max

	// Per-instance part of RenderInner (SWModelRenderer.cpp:49-178): origin shift, the 28 brightness levels, the sphere /
	// frustum cull, the screen-space basis and the point diameter.  One thread per instance.
	template <bool kSse>
	__global__ void k_model_setup(const SceneMats *__restrict__ mats, const InstanceIn *__restrict__ inst,
	                              InstanceSetup *__restrict__ out, int n, osbn::SseTables sse) {
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n)
			return;
		const osbn::Num<kSse> num{sse};
		const SceneMats &S = *mats;
		const InstanceIn &I = inst[i];
		const float *m = I.matrix;
		float origin[3] = {m[12], m[13], m[14]};
		const float a1[3] = {m[0], m[1], m[2]}, a2[3] = {m[4], m[5], m[6]}, a3[3] = {m[8], m[9], m[10]};
		madd3(origin, a1, __fadd_rn(I.model.origin[0], 0.25f));
		madd3(origin, a2, __fadd_rn(I.model.origin[1], 0.25f));
		madd3(origin, a3, __fadd_rn(I.model.origin[2], 0.25f));
		InstanceSetup R;
		{
			// brightness per quantised normal (:66-127).  As in the reference the innermost switch tests y, not z, and
			// starts again from the x term: the table depends on (x, y) only.
			const float lightVec[3] = {0.f, -0.707f, -0.707f};
			const float dot1 = __fmul_rn(dot3(a1, lightVec), num.rsqrt(len2_3(a1)));
			const float dot3v = __fmul_rn(dot3(a3, lightVec), num.rsqrt(len2_3(a3)));
			for (int x = 0; x < 3; x++) {
				const float d = x == 0 ? -dot1 : (x == 1 ? 0.f : dot1);
				const int cnt = x == 1 ? 0 : 1;
				for (int y = 0; y < 3; y++) {
					const int cnt2 = cnt + (y == 1 ? 0 : 1);
					for (int z = 0; z < 3; z++) {
						float d3 = d;
						int cnt3 = cnt2;
						if (y == 0) {
							d3 = __fsub_rn(d3, dot3v);
							cnt3++;
						} else if (y == 2) {
							d3 = __fadd_rn(d3, dot3v);
							cnt3++;
						}
						if (cnt3 == 2)
							d3 = __fmul_rn(d3, 0.707f);
						else if (cnt3 == 3)
							d3 = __fmul_rn(d3, 0.57735f);
						d3 = __fadd_rn(192.f, __fmul_rn(d3, 62.f));
						R.brights[x + y * 3 + z * 9] = (uint8_t)(osbn::x86_f2i(d3) & 0xff);
					}
				}
			}
			R.brights[27] = 255; // emissive
		}
		{
			float center[3] = {origin[0], origin[1], origin[2]};
			madd3(center, a1, I.model.center[0]);
			madd3(center, a2, I.model.center[1]);
			madd3(center, a3, I.model.center[2]);
			float largest = fmax_std(fmax_std(len2_3(a1), len2_3(a2)), len2_3(a3));
			const float radius = __fmul_rn(I.model.radius, __fsqrt_rn(largest));
			bool visible = true;
			for (int p = 0; p < 6; p++) { // SphereFrustrumCull (SWRenderer.cpp:1130-1147)
				const float dist = __fadd_rn(dot3(S.frustum[p], center), S.frustum[p][3]);
				if (dist < -radius)
					visible = false;
			}
			R.visible = visible ? 1 : 0;
		}
		const float sx = __fmul_rn((float)S.fw, 0.5f), sy = __fmul_rn((float)(-S.fh), 0.5f);
		float *dst[4] = {R.tOrigin, R.tAxis1, R.tAxis2, R.tAxis3};
		const float *src[4] = {origin, a1, a2, a3};
		float largest = 0.f;
		for (int k = 0; k < 4; k++) {
			mat_vec4(S.viewproj, src[k], k == 0 ? 1.f : 0.f, dst[k]);
			dst[k][0] = __fmul_rn(dst[k][0], sx);
			dst[k][1] = __fmul_rn(dst[k][1], sy);
			if (k > 0) {
				const float *t = dst[k];
				const float l = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(t[0], t[0]), __fmul_rn(t[1], t[1])), __fmul_rn(t[2], t[2])),
				                          __fmul_rn(t[3], t[3]));
				largest = k == 1 ? l : fmax_std(largest, l);
			}
		}
		R.pointDiameter = __fsqrt_rn(largest);
		out[i] = R;
	}

	__global__ void k_key_init(const float *__restrict__ depth, unsigned long long *__restrict__ keys, size_t n) {
		const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
		if (i < n)
			keys[i] = (unsigned long long)depth_key(depth[i]) << 32; // sequence 0: the terrain wins ties
	}

	// strict `zval < *db` on the packed key; the plain load first turns most overdraw into a read (a stale value is
	// only ever larger than the current one, so it can cause a redundant atomic, never a wrong skip)
#ifndef OSB_SPLAT_PRETEST
#define OSB_SPLAT_PRETEST 0 // 0: atomics only (fastest: the loads are latency, the atomics fire-and-forget), 1: L2 load first, 2: L1-cached load first
#endif
	__device__ __forceinline__ void key_min(unsigned long long *p, unsigned long long key) {
#if OSB_SPLAT_PRETEST == 1
		if (key < __ldcg(p))
#elif OSB_SPLAT_PRETEST == 2
		if (key < __ldca(p))
#endif
			atomicMin(p, key);
	}

	// The voxel loop of RenderInner (:184-276).  One WARP per (instance, column (x, y)): lane i takes the column's i-th
	// surface voxel (projection, cull, shading: the per-voxel float work runs 32 wide), then splats of up to
	// kOwnLaneArea pixels are written by their own lane and larger ones by the whole warp, one voxel at a time.
#ifndef OSB_SPLAT_OWN_AREA
#define OSB_SPLAT_OWN_AREA 16
#endif
	constexpr int kOwnLaneArea = OSB_SPLAT_OWN_AREA;
	template <bool kSse>
	__global__ void __launch_bounds__(128)
	  k_model_splat(const SceneMats *__restrict__ mats, const InstanceIn *__restrict__ inst,
	                const InstanceSetup *__restrict__ setup, int nInst, uint32_t totalCols,
	                unsigned long long *__restrict__ keys, uint32_t *__restrict__ splatColor, osbn::SseTables sse) {
		const uint32_t item = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
		const int lane = threadIdx.x & 31;
		if (item >= totalCols)
			return;
		int lo = 0, hi = nInst - 1;
		while (lo < hi) {
			const int mid = (lo + hi + 1) >> 1;
			if (inst[mid].colBase <= item)
				lo = mid;
			else
				hi = mid - 1;
		}
		const InstanceIn &I = inst[lo];
		const InstanceSetup &R = setup[lo];
		if (!R.visible)
			return;
		const osbn::Num<kSse> num{sse};
		const int fw = mats->fw, fh = mats->fh;
		const float zNear = mats->zNear;
		// traversal order is x outer, y inner (:185-188); item order inside the instance follows it
		const uint32_t col = item - I.colBase;
		const int x = (int)(col / (uint32_t)I.model.h), y = (int)(col % (uint32_t)I.model.h);
		const int ci = x + y * I.model.w, nCols = I.model.w * I.model.h;
		const int count = (int)I.model.addr[2 * nCols + ci]; // surface voxels of this column
		if (count == 0)
			return;
		float v2[4];
#pragma unroll
		for (int k = 0; k < 4; k++) {
			float v = R.tOrigin[k];
			for (int i = 0; i < x; i++)
				v = __fadd_rn(v, R.tAxis1[k]); // v1 += tAxis1
			for (int i = 0; i < y; i++)
				v = __fadd_rn(v, R.tAxis2[k]); // v2 += tAxis2
			v2[k] = v;
		}
		const uint32_t *mp = I.model.data + I.model.addr[ci];
		const uint32_t seq0 = I.seqBase + I.model.addr[nCols + ci]; // rank of the column's first voxel in traversal order
		const int offX = fw >> 1, offY = fh >> 1;
		for (int base = 0; base < count; base += 32) {
			const int vi = base + lane;
			bool have = vi < count;
			int minX = 0, minY = 0, rw = 0, rh = 0;
			unsigned long long key = 0;
			if (have) {
				const uint32_t data = mp[2 * vi], normal = mp[2 * vi + 1];
				const float zf = (float)(data >> 24);
				float vv[4];
#pragma unroll
				for (int k = 0; k < 4; k++)
					vv[k] = __fadd_rn(v2[k], __fmul_rn(R.tAxis3[k], zf));
				have = !(vv[2] < zNear);
				if (have) {
					const float scl = num.rcp(vv[3]);
					const int ix = osbn::x86_f2i(__fmul_rn(vv[0], scl)) + offX;
					const int iy = osbn::x86_f2i(__fmul_rn(vv[1], scl)) + offY;
					int idm = osbn::x86_f2i(__fadd_rn(__fmul_rn(R.pointDiameter, scl), .99f));
					idm = max(1, idm);
					minX = ix - (idm >> 1);
					minY = iy - (idm >> 1);
					int maxX = (int)((uint32_t)ix + (uint32_t)idm), maxY = (int)((uint32_t)iy + (uint32_t)idm);
					have = !(minX >= fw || minY >= fh) && !(maxX <= 0 || maxY <= 0);
					if (have) {
						minX = max(minX, 0);
						minY = max(minY, 0);
						rw = min(maxX, fw) - minX;
						rh = min(maxY, fh) - minY;
						uint32_t color = data & 0xffffffu;
						if (color == 0)
							color = I.customColor;
						const uint32_t bright = R.brights[normal < 28 ? normal : 27];
						color = ((((color & 0xff00u) * bright) & 0xff0000u) | (((color & 0xff00ffu) * bright) & 0xff00ff00u)) >> 8;
						const uint32_t seq = seq0 + (uint32_t)vi;
						splatColor[seq] = color;
						key = ((unsigned long long)depth_key(vv[2]) << 32) | seq;
					}
				}
			}
			const bool own = have && rw * rh <= kOwnLaneArea;
			if (own) {
				unsigned long long *row = keys + (size_t)minY * fw + minX;
				for (int yy = 0; yy < rh; yy++, row += fw)
					for (int xx = 0; xx < rw; xx++)
						key_min(row + xx, key);
			}
			unsigned big = __ballot_sync(0xffffffffu, have && !own);
			while (big) {
				const int src = __ffs(big) - 1;
				big &= big - 1;
				const int bx = __shfl_sync(0xffffffffu, minX, src), by = __shfl_sync(0xffffffffu, minY, src);
				const int bw = __shfl_sync(0xffffffffu, rw, src), bh = __shfl_sync(0xffffffffu, rh, src);
				const unsigned long long bk = __shfl_sync(0xffffffffu, key, src);
				unsigned long long *org = keys + (size_t)by * fw + bx;
				if (bw <= 32) { // lanes tile the rectangle: 2^sh columns x (32 >> sh) rows per pass
					const int sh = bw <= 1 ? 0 : 32 - __clz(bw - 1);
					const int lx = lane & ((1 << sh) - 1), ly = lane >> sh, stepY = 32 >> sh;
					if (lx < bw)
						for (int yy = ly; yy < bh; yy += stepY)
							key_min(org + (size_t)yy * fw + lx, bk);
				} else {
					for (int yy = 0; yy < bh; yy++, org += fw)
						for (int xx = lane; xx < bw; xx += 32)
							key_min(org + xx, bk);
				}
			}
		}
	}

	// The float recurrences of ApplyDynamicLight (:266-272,327-333): vx restarts at the light's minX and grows by dvx per
	// pixel; vy restarts at the first row of every emulated worker thread's band and grows by dvy per row.
	__global__ void k_light_prep(const LightDev *__restrict__ lights, int nLights, int fw, int fh, int swThreads, float fovX,
	                             float fovY, float dvx, float dvy, float *__restrict__ tabs) {
		const int li = blockIdx.x;
		if (li >= nLights)
			return;
		const LightDev &L = lights[li];
		float *vxT = tabs + L.tabBase, *vyT = vxT + fw;
		if (threadIdx.x == 0) {
			float vx = __fadd_rn(fovX, __fmul_rn(dvx, (float)L.minX));
			for (int x = L.minX; x < L.maxX; x++) {
				vxT[x] = vx;
				vx = __fadd_rn(vx, dvx);
			}
		}
		const unsigned lightHeight = (unsigned)(L.maxY - L.minY);
		for (unsigned t = threadIdx.x; t < (unsigned)swThreads; t += blockDim.x) {
			const int startY = (int)(lightHeight * t / (unsigned)swThreads) + L.minY;
			const int endY = (int)(lightHeight * (t + 1) / (unsigned)swThreads) + L.minY;
			float vy = __fadd_rn(fovY, __fmul_rn(dvy, (float)startY));
			for (int y = startY; y < endY; y++) {
				vyT[y] = vy;
				vy = __fadd_rn(vy, dvy);
			}
		}
	}

	template <class T> static int grow_buf(T *&p, size_t &cap, size_t want, bool pinned = false) {
		if (want <= cap)
			return 0;
		if (p) {
			if (pinned)
				cudaFreeHost(p);
			else
				cudaFree(p);
		}
		p = nullptr;
		cap = 0;
		const size_t n = want + (want >> 1) + 16;
		if (pinned)
			OSB_CUDA(cudaMallocHost((void **)&p, n * sizeof(T)));
		else
			OSB_CUDA(cudaMalloc((void **)&p, n * sizeof(T)));
		cap = n;
		return 0;
	}

	static int models_init(osb200_ctx *c) {
		if (c->models)
			return 0;
		auto *M = new ModelState();
		c->models = M;
		OSB_CUDA(cudaMalloc(&M->d_mats, sizeof(SceneMats)));
		OSB_CUDA(cudaMallocHost(&M->h_mats, sizeof(SceneMats)));
		OSB_CUDA(cudaEventCreateWithFlags(&M->stageFree, cudaEventDisableTiming));
		OSB_CUDA(cudaStreamCreateWithFlags(&M->aux, cudaStreamNonBlocking));
		OSB_CUDA(cudaEventCreateWithFlags(&M->auxDone, cudaEventDisableTiming));
		return 0;
	}

	void models_destroy(osb200_ctx *c) {
		ModelState *M = c->models;
		if (!M)
			return;
		cudaFree(M->d_inst);
		cudaFreeHost(M->h_inst);
		cudaFree(M->d_setup);
		cudaFree(M->d_mats);
		cudaFreeHost(M->h_mats);
		cudaFree(M->d_keys);
		cudaFree(M->d_splatColor);
		cudaFree(M->d_lights);
		cudaFreeHost(M->h_lights);
		cudaFree(M->d_lightTabs);
		if (M->stageFree)
			cudaEventDestroy(M->stageFree);
		if (M->auxDone)
			cudaEventDestroy(M->auxDone);
		if (M->aux)
			cudaStreamDestroy(M->aux);
		delete M;
		c->models = nullptr;
	}

	// ---- host float helpers in the reference's operation order ----
	static inline float hdot3(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

	// SceneDefinition::ToOpenGLProjectionMatrix / ToViewMatrix, Matrix4Multiply, SWRenderer::BuildFrustrum
	// (SceneDefinition.cpp:26-72, Math.cpp:108-128, SWRenderer.cpp:207-227)
	static void fill_scene_mats(const osb200_scene &s, int fw, int fh, SceneMats &S) {
		const float near = s.zNear, far = s.zFar;
		const float t = near * std::tan(s.fovY * .5f), r = near * std::tan(s.fovX * .5f);
		const float a = r * 2.f, b = t * 2.f, cc = far - near;
		float P[16] = {0}, V[16] = {0};
		P[0] = near * 2.f / a;
		P[5] = near * 2.f / b;
		P[10] = -(far + near) / cc;
		P[11] = -1.f;
		P[14] = -(far * near * 2.f) / cc;
		V[15] = 1.f;
		for (int k = 0; k < 3; k++) {
			V[k * 4 + 0] = s.axis[0][k];
			V[k * 4 + 1] = s.axis[1][k];
			V[k * 4 + 2] = -s.axis[2][k];
		}
		{ // v = mat * viewOrigin with a zero translation column, then the negated result becomes the translation
			float v[3];
			for (int rr = 0; rr < 3; rr++) {
				float acc = V[rr] * s.origin[0];
				acc += V[4 + rr] * s.origin[1];
				acc += V[8 + rr] * s.origin[2];
				acc += V[12 + rr] * 1.f;
				v[rr] = acc;
			}
			V[12] = -v[0];
			V[13] = -v[1];
			V[14] = -v[2];
		}
		// projectionMatrix * viewMatrix: out[4c + r] = b[4c]*a[r] + b[4c+1]*a[4+r] + b[4c+2]*a[8+r] + b[4c+3]*a[12+r]
		for (int col = 0; col < 4; col++)
			for (int row = 0; row < 4; row++)
				S.viewproj[col * 4 + row] = V[col * 4] * P[row] + V[col * 4 + 1] * P[4 + row] + V[col * 4 + 2] * P[8 + row] +
				                            V[col * 4 + 3] * P[12 + row];
		auto plane = [&](int i, const float n[3]) { // Plane3::PlaneWithPointOnPlane
			for (int k = 0; k < 3; k++)
				S.frustum[i][k] = n[k];
			S.frustum[i][3] = -hdot3(s.origin, n);
		};
		plane(0, s.axis[2]);
		for (int k = 0; k < 4; k++)
			S.frustum[1][k] = -S.frustum[0][k]; // Flipped()
		S.frustum[0][3] -= s.zNear;
		S.frustum[1][3] += s.zFar;
		const float xCos = cosf(s.fovX * .5f), xSin = sinf(s.fovX * .5f), yCos = cosf(s.fovY * .5f), ySin = sinf(s.fovY * .5f);
		float n[3];
		for (int k = 0; k < 3; k++)
			n[k] = s.axis[2][k] * xSin - s.axis[0][k] * xCos;
		plane(2, n);
		for (int k = 0; k < 3; k++)
			n[k] = s.axis[2][k] * xSin + s.axis[0][k] * xCos;
		plane(3, n);
		for (int k = 0; k < 3; k++)
			n[k] = s.axis[2][k] * ySin - s.axis[1][k] * yCos;
		plane(4, n);
		for (int k = 0; k < 3; k++)
			n[k] = s.axis[2][k] * ySin + s.axis[1][k] * yCos;
		plane(5, n);
		S.zNear = s.zNear;
		S.fw = fw;
		S.fh = fh;
	}

	// SWRenderer::AddLight (SWRenderer.cpp:578-649): fog cull, view-plane cull and the screen rectangle of a point light.
	// Returns false when the light is dropped.
	static bool add_light(const osb200_scene &s, const osb200_light &p, int fw, int fh, float fogDistance, LightDev &L) {
		float diff[3];
		for (int k = 0; k < 3; k++)
			diff[k] = p.origin[k] - s.origin[k];
		const float rad2 = p.radius * p.radius;
		const float poweredLength = hdot3(diff, diff);
		const float fogCullRange = p.radius + fogDistance;
		if (poweredLength > fogCullRange * fogCullRange)
			return false;
		if (poweredLength < rad2) {
			L.minX = 0;
			L.minY = 0;
			L.maxX = fw;
			L.maxY = fh;
		} else if (hdot3(diff, s.axis[2]) < 0.f) {
			return false;
		} else {
			auto viewRange = [](float cx, float cy, float fov, float screenSize, int out[2]) {
				auto trans = [screenSize, fov](float v) {
					v = (v / fov) * 0.5f + 0.5f;
					v = std::max(v, 0.f);
					v = std::min(v, 1.f);
					v *= screenSize;
					return static_cast<int>(v);
				};
				float dist = cx * cx + cy * cy - 1.f;
				if (dist <= 0.f) {
					out[0] = 0;
					out[1] = static_cast<int>(screenSize);
					return;
				}
				float denom = cx * cx - 1.f;
				if (fabsf(denom) < 1.e-10f)
					denom = 1.e-8f;
				denom = 1.f / denom; // computed and never used by the reference either
				(void)denom;
				dist = sqrtf(dist);
				if (cx <= 1.f) {
					if (cy > 0.f) {
						out[0] = trans(cx * cy - dist);
						out[1] = static_cast<int>(screenSize);
					} else {
						out[0] = 0;
						out[1] = trans(cx * cy + dist);
					}
				} else {
					out[0] = trans(cx * cy - dist);
					out[1] = trans(cx * cy + dist);
				}
			};
			const float invRad = 1.f / p.radius;
			int rx[2], ry[2];
			viewRange(hdot3(diff, s.axis[2]) * invRad, hdot3(diff, s.axis[0]) * invRad, tanf(s.fovX * 0.5f), (float)fw, rx);
			viewRange(hdot3(diff, s.axis[1]) * invRad, hdot3(diff, s.axis[0]) * invRad, tanf(s.fovY * 0.5f), (float)fh, ry);
			L.minX = rx[0];
			L.maxX = rx[1];
			L.minY = ry[0];
			L.maxY = ry[1];
		}
		// ApplyDynamicLight's per-light scalars (:252-263)
		L.center[0] = hdot3(diff, s.axis[0]);
		L.center[1] = hdot3(diff, s.axis[1]);
		L.center[2] = hdot3(diff, s.axis[2]);
		auto toFactor8 = [](float v) { // ToFixedFactor8, SWUtils.h:73-76
			int i = static_cast<int>(v * 256.f + .5f);
			return std::max(std::min(i, 256), 0);
		};
		L.r = toFactor8(p.color[0]);
		L.g = toFactor8(p.color[1]);
		L.b = toFactor8(p.color[2]);
		L.invRadius2 = 1.f / (p.radius * p.radius);
		return true;
	}
} // namespace osb

extern "C" {

int osb200_model_create(osb200_ctx *c, int w, int h, int d, const float origin[3], const uint64_t *solid,
                        const uint32_t *colors, osb200_model **out) {
	if (!c || !origin || !solid || !colors || !out || w <= 0 || h <= 0 || d <= 0 || d > 64 || w > 4096 || h > 4096) {
		osb::set_error("osb200_model_create: bad argument (1 <= depth <= 64)");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	auto *M = new osb200_model();
	M->w = w;
	M->h = h;
	M->d = d;
	M->owner = c;
	M->device = c->device;
	for (int k = 0; k < 3; k++)
		M->origin[k] = origin[k];
	// SWModel.cpp:28-32: center = (w, h, d) * 0.5, radius = |center|
	M->center[0] = (float)w * 0.5f;
	M->center[1] = (float)h * 0.5f;
	M->center[2] = (float)d * 0.5f;
	M->radius = sqrtf(M->center[0] * M->center[0] + M->center[1] * M->center[1] + M->center[2] * M->center[2]);
	osb::build_render_data(*M, solid, colors);
	// behind the column start table: the rank of each column's first voxel in the renderer's x-outer / y-inner order
	std::vector<uint32_t> addrAndRank(M->renderDataAddr);
	addrAndRank.resize((size_t)w * h * 3); // + the number of surface voxels of each column
	uint32_t rank = 0;
	for (int x = 0; x < w; x++)
		for (int y = 0; y < h; y++) {
			addrAndRank[(size_t)w * h + x + y * w] = rank;
			const uint32_t a = M->renderDataAddr[x + y * w];
			const uint32_t e = (x + y * w + 1 < w * h) ? M->renderDataAddr[x + y * w + 1] : (uint32_t)M->renderData.size();
			addrAndRank[(size_t)w * h * 2 + x + y * w] = (e - a - 1) / 2;
			rank += (e - a - 1) / 2;
		}
	cudaError_t e1 = cudaMalloc(&M->d_data, M->renderData.size() * 4);
	cudaError_t e2 = cudaMalloc(&M->d_addr, addrAndRank.size() * 4);
	if (e1 != cudaSuccess || e2 != cudaSuccess) {
		cudaFree(M->d_data);
		cudaFree(M->d_addr);
		delete M;
		osb::set_error("osb200_model_create: cudaMalloc failed");
		return -2;
	}
	cudaMemcpy(M->d_data, M->renderData.data(), M->renderData.size() * 4, cudaMemcpyHostToDevice);
	cudaMemcpy(M->d_addr, addrAndRank.data(), addrAndRank.size() * 4, cudaMemcpyHostToDevice);
	*out = M;
	return 0;
}

void osb200_model_destroy(osb200_model *m) {
	if (!m)
		return;
	cudaSetDevice(m->device);
	cudaDeviceSynchronize();
	cudaFree(m->d_data);
	cudaFree(m->d_addr);
	delete m;
}

int osb200_model_render_data(const osb200_model *m, uint32_t *data, int dataCap, uint32_t *addr, int addrCap) {
	if (!m)
		return -1;
	if (data && dataCap >= (int)m->renderData.size())
		std::memcpy(data, m->renderData.data(), m->renderData.size() * 4);
	if (addr && addrCap >= (int)m->renderDataAddr.size())
		std::memcpy(addr, m->renderDataAddr.data(), m->renderDataAddr.size() * 4);
	return (int)m->renderData.size();
}

int osb200_render_scene(osb200_ctx *c, const osb200_scene *scene, int w, int h, const float fogColor[3], float fogDist,
                        const osb200_model_instance *inst, int nInst, const osb200_light *lights, int nLights,
                        uint32_t *host_bgrx, float *host_depth) {
	using namespace osb;
	if (!c || !scene || !fogColor || nInst < 0 || nLights < 0 || (nInst && !inst) || (nLights && !lights)) {
		set_error("osb200_render_scene: bad argument");
		return -1;
	}
	if (w <= 0 || h <= 0 || (w & 7) || (h & 7)) {
		set_error("Framebuffer size is not multiple of 8."); // SWRenderer.cpp:105-107
		return -1;
	}
	for (int i = 0; i < nInst; i++)
		if (!inst[i].model || inst[i].model->owner != c) {
			set_error("osb200_render_scene: instance %d has no model of this context", i);
			return -1;
		}
	OSB_CUDA(cudaSetDevice(c->device));
	if (models_init(c))
		return -2;
	ModelState *M = c->models;
	uint32_t *d_fb;
	float *d_depth;
	if (int rc = terrain_out_buffers(c, (size_t)w * h, &d_fb, &d_depth))
		return rc;

	// lights that survive AddLight's culls
	std::vector<LightDev> kept;
	for (int i = 0; i < nLights; i++) {
		LightDev L;
		if (add_light(*scene, lights[i], w, h, fogDist, L)) {
			L.tabBase = (uint32_t)(kept.size() * (size_t)(w + h));
			kept.push_back(L);
		}
	}
	const bool extras = nInst > 0 || !kept.empty();
	// terrain (+ fog right away when nothing is drawn between them)
	if (int rc = render_impl(c, scene, 1, w, h, fogColor, fogDist, d_fb, d_depth, nullptr, nullptr, extras))
		return rc;
	const size_t npix = (size_t)w * h;
	osbn::SseTables sse{c->d_rcp, c->d_rsqrt, c->rcpBits, c->rsqrtBits};
	const bool kSseMode = c->opts.numerics != OSB200_NUM_IEEE;
	OSB_CUDA(cudaEventSynchronize(M->stageFree)); // pinned staging of the previous call
	// The terrain kernels are already queued on the context stream; the per-instance set-up and the light tables do not
	// depend on them and go to a second stream (with per-kernel timing on, everything stays on the context stream).
	cudaStream_t aux = c->timing ? c->stream : M->aux;
	uint32_t totalCols = 0;
	if (nInst > 0) {
		if (grow_buf(M->h_inst, M->instCap, (size_t)nInst, true))
			return -2;
		if (grow_buf(M->d_inst, M->dInstCap, (size_t)nInst) || grow_buf(M->d_setup, M->dSetupCap, (size_t)nInst))
			return -2;
		auto toFixed8 = [](float v) { // SWUtils.h:68-71
			int i = static_cast<int>(v * 255.f + .5f);
			return (uint32_t)std::max(std::min(i, 255), 0);
		};
		// The heaviest instances (estimated covered pixels = voxels x (scale / distance)^2) are launched first so that
		// a near, large-splat instance does not end up as the tail of the splat kernel.  Only the launch order
		// changes: sequence numbers keep the submission order, which is what decides ties.
		std::vector<int> order(nInst);
		std::vector<float> cost(nInst);
		std::vector<uint32_t> seqBase(nInst);
		uint32_t seq = 1, cols = 0;
		for (int i = 0; i < nInst; i++) {
			const float *m = inst[i].matrix;
			float s2 = 0.f, d2 = 0.f;
			for (int a = 0; a < 3; a++)
				s2 = std::max(s2, m[a * 4] * m[a * 4] + m[a * 4 + 1] * m[a * 4 + 1] + m[a * 4 + 2] * m[a * 4 + 2]);
			for (int k = 0; k < 3; k++)
				d2 += (m[12 + k] - scene->origin[k]) * (m[12 + k] - scene->origin[k]);
			cost[i] = (float)inst[i].model->voxels * s2 / std::max(d2, 1.f);
			order[i] = i;
			seqBase[i] = seq;
			seq += inst[i].model->voxels;
		}
		{ // the kHeavyFirst most expensive instances go to the front, the rest keep the submission order (sorting all
			// of them makes the near, mutually overlapping instances run at the same time and contend in L2: measured
			// +8 % at 2048 instances, while the front group alone removes the tail: -18 % at 64)
			constexpr int kHeavyFirst = 16;
			std::vector<int> byCost(order);
			const int nh = std::min(kHeavyFirst, nInst);
			std::partial_sort(byCost.begin(), byCost.begin() + nh, byCost.end(),
			                  [&](int a, int b) { return cost[a] > cost[b] || (cost[a] == cost[b] && a < b); });
			std::vector<char> front(nInst, 0);
			int slot = 0;
			for (int k = 0; k < nh; k++) {
				order[slot++] = byCost[k];
				front[byCost[k]] = 1;
			}
			for (int i = 0; i < nInst; i++)
				if (!front[i])
					order[slot++] = i;
		}
		for (int slot = 0; slot < nInst; slot++) {
			const int i = order[slot];
			const osb200_model *mo = inst[i].model;
			InstanceIn &I = M->h_inst[slot];
			I.model.data = mo->d_data;
			I.model.addr = mo->d_addr;
			I.model.w = mo->w;
			I.model.h = mo->h;
			for (int k = 0; k < 3; k++) {
				I.model.origin[k] = mo->origin[k];
				I.model.center[k] = mo->center[k];
			}
			I.model.radius = mo->radius;
			std::memcpy(I.matrix, inst[i].matrix, sizeof(I.matrix));
			I.customColor = toFixed8(inst[i].customColor[2]) | (toFixed8(inst[i].customColor[1]) << 8) |
			                (toFixed8(inst[i].customColor[0]) << 16);
			I.seqBase = seqBase[i];
			I.colBase = cols;
			cols += (uint32_t)(mo->w * mo->h);
		}
		fill_scene_mats(*scene, w, h, *M->h_mats);
		if (grow_buf(M->d_keys, M->keyCap, npix) || grow_buf(M->d_splatColor, M->splatCap, (size_t)seq))
			return -2;
		OSB_CUDA(cudaMemcpyAsync(M->d_mats, M->h_mats, sizeof(SceneMats), cudaMemcpyHostToDevice, aux));
		OSB_CUDA(cudaMemcpyAsync(M->d_inst, M->h_inst, sizeof(InstanceIn) * nInst, cudaMemcpyHostToDevice, aux));
		{
			KLaunch kl(c, K_MODEL_SETUP, aux);
			if (kSseMode)
				k_model_setup<true><<<(nInst + 63) / 64, 64, 0, aux>>>(M->d_mats, M->d_inst, M->d_setup, nInst, sse);
			else
				k_model_setup<false><<<(nInst + 63) / 64, 64, 0, aux>>>(M->d_mats, M->d_inst, M->d_setup, nInst, sse);
		}
		totalCols = cols;
	}
	if (!kept.empty()) {
		const size_t nl = kept.size();
		if (grow_buf(M->h_lights, M->lightCap, nl, true))
			return -2;
		if (grow_buf(M->d_lights, M->dLightCap, nl))
			return -2;
		if (grow_buf(M->d_lightTabs, M->lightTabCap, nl * (size_t)(w + h)))
			return -2;
		std::memcpy(M->h_lights, kept.data(), nl * sizeof(LightDev));
		OSB_CUDA(cudaMemcpyAsync(M->d_lights, M->h_lights, nl * sizeof(LightDev), cudaMemcpyHostToDevice, aux));
		const float fovX = tanf(scene->fovX * 0.5f), fovY = tanf(scene->fovY * 0.5f);
		const float dvx = -fovX * 2.f / static_cast<float>(w), dvy = -fovY * 2.f / static_cast<float>(h);
		KLaunch kl(c, K_LIGHTS, aux);
		k_light_prep<<<(unsigned)nl, 32, 0, aux>>>(M->d_lights, (int)nl, w, h, c->opts.sw_threads, fovX, fovY, dvx, dvy,
		                                        M->d_lightTabs);
	}
	if (aux != c->stream) { // join: what follows reads the set-up records / light tables and the terrain's depth
		OSB_CUDA(cudaEventRecord(M->auxDone, aux));
		OSB_CUDA(cudaStreamWaitEvent(c->stream, M->auxDone, 0));
	}
	if (nInst > 0) {
		KLaunch kl(c, K_MODEL_SPLAT);
		k_key_init<<<(unsigned)((npix + 255) / 256), 256, 0, c->stream>>>(d_depth, M->d_keys, npix);
		if (kSseMode)
			k_model_splat<true><<<(totalCols + 3) / 4, 128, 0, c->stream>>>(M->d_mats, M->d_inst, M->d_setup, nInst, totalCols,
			                                                                    M->d_keys, M->d_splatColor, sse);
		else
			k_model_splat<false><<<(totalCols + 3) / 4, 128, 0, c->stream>>>(M->d_mats, M->d_inst, M->d_setup, nInst, totalCols,
			                                                                     M->d_keys, M->d_splatColor, sse);
		c->launches++; // k_key_init rides in the splat bracket; the winners are resolved by the finishing pass below
	}
	OSB_CUDA(cudaEventRecord(M->stageFree, c->stream));
	// splat winners -> lights in submission order -> fog, one pass over the frame (k_scene_finish, terrain.cu)
	if (extras)
		if (int rc = terrain_finish_scene(c, w, h, fogColor, d_fb, d_depth, nInst > 0 ? M->d_keys : nullptr, M->d_splatColor,
		                                  M->d_lights, (int)kept.size(), M->d_lightTabs))
			return rc;
	OSB_CUDA(cudaGetLastError());
	if (host_bgrx)
		OSB_CUDA(cudaMemcpyAsync(host_bgrx, d_fb, npix * 4, cudaMemcpyDeviceToHost, c->stream));
	if (host_depth)
		OSB_CUDA(cudaMemcpyAsync(host_depth, d_depth, npix * 4, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

} // extern "C"
