// CPU lock-step emulation of the terrain kernels (TEST INFRASTRUCTURE, never part of the product).
//
// openspades_b200/csrc/terrain_core.cuh holds the per-lane code of k_line_setup, k_walk_lines and k_gather_fog; this
// file compiles that very header with g++ and runs the lanes one after the other in the kernels' launch structure
// (per-view set-up, pitch cull + compaction of the surviving lines, one LineWalker per active line, one GatherLane x 32
// per warp task with the corner / pixel phases in kernel order).  tests/test_terrain_emul.py compares the frames with
// the oracle, so control-flow and arithmetic mistakes in the core are found in the CPU container (no GPU here).
//
// g++ -O2 -ffp-contract=off -shared -fPIC terrain_emul.cpp ../../openspades_b200/csrc/sse_calibrate.cpp
#include <cstdarg>
#include <cstdio>
#include <algorithm>
#include <vector>

// per-lane counters of the current step (span candidates, landing words, stored pixels): the warp model below turns them
// into what a 32-lane warp would execute
static thread_local unsigned g_stat[4];
#define OSB_STAT(what, n) (g_stat[what] += (n))
#include "../../openspades_b200/csrc/terrain_core.cuh"

namespace osb {
	void set_error(const char *fmt, ...) {
		va_list ap;
		va_start(ap, fmt);
		vfprintf(stderr, fmt, ap);
		va_end(ap);
		fputc('\n', stderr);
	}
} // namespace osb

extern "C" int osb200_calibrate_sse(uint32_t *rcp_table, int *rcp_bits, uint32_t *rsqrt_table, int *rsqrt_bits, int max_bits);

namespace {
	using namespace osb;
	unsigned long long g_coopStats[4]; // lines, fallbacks, max spans, sum spans
	int g_coop = 0; // 0: lane-per-line walker, 1: cooperative walker (one warp per line), 2: its sequential fallback

	struct Tables {
		std::vector<uint32_t> rcp, rsqrt;
		int rcpBits = 0, rsqrtBits = 0;
		bool ok = false;
	};
	Tables &tables() {
		static Tables t = [] {
			Tables x;
			x.rcp.resize(1 << 16);
			x.rsqrt.resize(2 << 16);
			x.ok = osb200_calibrate_sse(x.rcp.data(), &x.rcpBits, x.rsqrt.data(), &x.rsqrtBits, 16) == 0;
			return x;
		}();
		return t;
	}

	std::vector<uint16_t> atan_table() { // AtanTable (SWMapRenderer.cpp:673-696), as terrain.cu builds it
		std::vector<uint16_t> tab(20000);
		auto toFixed = [](float v) -> uint16_t {
			v /= (M_PI * 2.f);
			v *= 65536.f;
			int i = static_cast<int>(v);
			return static_cast<uint16_t>(i & 65535);
		};
		for (int i = 0; i < 5000; i++) {
			tab[i] = toFixed(atanf(i / 4096.f));
			tab[5000 + i] = toFixed(atanf(1.f / ((i + .5f) / 4096.f)));
			tab[10000 + i] = toFixed(-atanf(i / 4096.f));
			tab[15000 + i] = toFixed(-atanf(1.f / ((i + .5f) / 4096.f)));
		}
		return tab;
	}

	struct Stats {
		unsigned long long lines, active, steps;
	};
	template <bool kSse, bool kSse2> void warp_model(WalkIO wio, const osbn::Num<kSse> &num, const std::vector<ViewParams> &views,
	                                                 const std::vector<LineRec> &recs, const std::vector<uint32_t> &active,
	                                                 unsigned long long *out);

	template <bool kSse, bool kSse2> int run(const uint64_t *solidT, const uint32_t *colorT, const osb200_scene *scenes, int nViews, int w,
	                                         int h, const float fog[3], float fogDist, int threads, int under, int noMap, int noFog,
	                                         uint32_t *fb, float *depth, Stats *stats, unsigned long long *modelOut = nullptr) {
		Tables &T = tables();
		if (kSse && !T.ok)
			return -4;
		osbn::SseTables sse{T.rcp.data(), T.rsqrt.data(), T.rcpBits, T.rsqrtBits};
		osbn::Num<kSse> num{sse};
		FrameGeom g{};
		g.fw = w;
		g.fh = h;
		g.under = under;
		g.swThreads = threads;
		g.sse2Fog = g.sse2Shade = kSse2;
		g.noMap = noMap;
		g.noFog = noFog;
		auto toFixed8 = [](float v) {
			int i = static_cast<int>(v * 255.f + .5f);
			return (uint32_t)std::max(std::min(i, 255), 0);
		};
		g.fogR = toFixed8(fog[0]);
		g.fogG = toFixed8(fog[1]);
		g.fogB = toFixed8(fog[2]);

		// ---- per-view set-up (host half + k_view_prep) ----
		std::vector<ViewParams> views(nViews);
		const uint32_t tabF = tab_floats(w, h);
		size_t lines = 0;
		for (int i = 0; i < nViews; i++) {
			fill_view(scenes[i], w, h, under, fogDist, views[i]);
			views[i].lineBase = (uint32_t)lines;
			views[i].tabBase = (uint32_t)i * tabF;
			lines += views[i].numLines;
		}
		const int lineRes = views[0].lineRes, words = lineRes >> 5;
		std::vector<float2> lineDirs(lines), lineTan(lines);
		std::vector<float> tabs((size_t)nViews * tabF);
		for (int i = 0; i < nViews; i++) {
			const ViewParams &V = views[i];
			float *vx = tabs.data() + V.tabBase, *vy = vx + (w >> 2), *v1 = vy + (h >> 2);
			prep_line_dirs(V, lineDirs.data() + V.lineBase);
			prep_fog_vx(V, g, vx);
			for (int t = 0; t < threads; t++) {
				prep_fog_vy(V, g, t, vy);
				prep_v1(V, g, t, v1);
			}
		}
		// ---- k_line_setup: pitch cull, records, compaction ----
		std::vector<LineRec> recs(lines);
		std::vector<uint32_t> active;
		if (!noMap) {
			for (int i = 0; i < nViews; i++) {
				const ViewParams &V = views[i];
				for (int l = 0; l < V.numLines; l++) {
					const uint32_t gl = V.lineBase + l;
					const LineSetupOut o = line_setup<kSse>(V, lineDirs[gl], lineRes, num);
					lineTan[gl] = make_float2(o.minTan, o.pitchScale);
					int slot = -1;
					if (!o.culled) {
						slot = (int)active.size();
						active.push_back(gl);
					}
					recs[gl] = make_line_rec(o, slot, (OSB_STRIP_HULL && slot < 0) ? kEmptyHull : i, lineRes);
				}
			}
		} else {
			for (size_t gl = 0; gl < lines; gl++)
				recs[gl] = LineRec{0, 0, -1, kEmptyHull};
		}
		// strips are deliberately filled with garbage: the walker never clears them, the gather must not read an
		// unwritten pixel
		std::vector<uint2> strips(active.size() * (size_t)lineRes, make_uint2(0xdeadbeefu, 0x7fc00000u));
		std::vector<uint32_t> mask(active.size() * (size_t)words, 0xffffffffu);
		// ---- k_walk_lines ----
		WalkIO wio{views.data(), lineDirs.data(), lineTan.data(), recs.data(), solidT, colorT, strips.data(), mask.data(), lineRes};
		if (modelOut) { // lock-step warp statistics instead of a frame (on copies: the walk rewrites the records)
			std::vector<LineRec> recs2 = recs;
			WalkIO wio2 = wio;
			wio2.lineRecs = recs2.data();
			warp_model<kSse, kSse2>(wio2, num, views, recs, active, modelOut);
			return 0;
		}
		unsigned long long steps = 0;
		if (g_coop) { // k_walk_coop: one warp per line, the phases in kernel order, the lanes one after the other
			std::vector<uint32_t> shbuf((coop_shared_bytes(lineRes) + 3) / 4);
			CoopShared &sh = *reinterpret_cast<CoopShared *>(shbuf.data());
			uint32_t *key = shbuf.data() + sizeof(CoopShared) / 4;
			for (size_t s = 0; s < active.size(); s++) {
				CoopLine<kSse, kSse2> C[32];
				for (int l = 0; l < 32; l++)
					C[l].begin(wio, active[s], (uint32_t)s, sh, key, l);
				C[0].path(sh);
				steps += sh.nSteps;
				bool ended = false;
				for (int s0 = 0; s0 < sh.nSteps && !ended; s0 += 32) {
					int cullAt = 1 << 30;
					for (int l = 31; l >= 0; l--)
						if (C[l].culls_after(wio, num, sh, s0 + l))
							cullAt = s0 + l;
					for (int l = 31; l >= 0; l--) { // any lane order must give the same strip: run them backwards
						const int st = s0 + l;
						if (st < sh.nSteps && st <= cullAt)
							C[l].emit_step(wio, num, sh, key, st);
					}
					ended = cullAt < (1 << 30);
				}
				g_coopStats[0]++;
				g_coopStats[2] = std::max<unsigned long long>(g_coopStats[2], (unsigned long long)sh.nSpans);
				g_coopStats[3] += sh.nSpans;
				if (sh.nSpans > kCoopMaxSpans)
					g_coopStats[1]++;
				if (sh.nSpans > kCoopMaxSpans || g_coop == 2) { // fallback: the sequential walker (g_coop == 2 forces it)
					LineWalker<kSse, kSse2> W;
					W.init(wio, active[s], (uint32_t)s, key, 1);
					while (W.step(wio, num, key, 1)) {
					}
					continue;
				}
				for (int l = 0; l < 32; l++)
					C[l].resolve_long(sh, key, l);
				int first = lineRes, last = -1;
				for (int l = 0; l < 32; l++) {
					int f, la;
					C[l].scan_hull(wio, key, l, f, la);
					first = std::min(first, f);
					last = std::max(last, la);
				}
				for (int l = 0; l < 32; l++)
					C[l].store(wio, sh, key, l, first, last);
			}
		} else
		for (size_t s = 0; s < active.size(); s++) {
			uint32_t filled[128 + kWalkExtraWords];
			LineWalker<kSse, kSse2> W;
			W.init(wio, active[s], (uint32_t)s, filled, 1);
			while (W.step(wio, num, filled, 1))
				steps++;
		}
		if (stats) {
			stats->lines = lines;
			stats->active = active.size();
			stats->steps = steps;
		}
		// ---- k_gather_fog ----
		const int segRowsArg = (h >> 3) % 3 == 0 ? 2 : ((w >> 3) & 1 ? 4 : kSegRows); // exercise all task heights across the test sizes
		const std::vector<uint16_t> atanTab = atan_table();
		GatherIO gio{views.data(), nViews,       g,         tabs.data(), solidT, atanTab.data(), strips.data(), recs.data(),
		             lineTan.data(), mask.data(), lineRes, fb,          depth, segRowsArg};
		const int colGroups = ((w >> 3) + 3) >> 2, segs = ((h >> 3) + segRowsArg - 1) / segRowsArg;
		const uint32_t tasks = (uint32_t)nViews * colGroups * segs;
		for (uint32_t task = 0; task < tasks; task++) {
			GatherLane<kSse, kSse2> L[32];
			GatherShared sm;
			for (int l = 0; l < 32; l++)
				if (!L[l].init(gio, task, l))
					return -9;
			for (int l = 0; l < 32; l++)
				L[l].fog_phase(gio, num, sm, l);
			for (int by = L[0].by0; by < L[0].by1; by += 2) {
				for (int l = 0; l < 32; l++)
					L[l].corner_phase(gio, num, sm, l, by);
				for (int r = 0; r < 2; r++)
					for (int l = 0; l < 32; l++)
						L[l].pixel_phase(gio, num, sm, r, by);
			}
		}
		return 0;
	}
} // namespace

static std::vector<float> g_taskCost; // per walker task: modelled warp instructions (see warp_model)
static std::vector<int> g_taskView;
static std::vector<int> g_taskSteps, g_taskPred; // per task: max DDA steps of a line, and their prediction at set-up
namespace {
	// Lock-step model of k_walk_lines' warps on one view batch: tasks of 32 adjacent surviving lines of a view, all lanes
	// step together until the last one ends.  Reports where lanes idle: out[0] warp-steps, [1] lane-steps, [2] sum over
	// warp-steps of the MAX span candidates of a lane (= iterations of the span loop), [3] sum of candidates, [4] warp-level
	// fill events (iterations with >= 1 landing lane) approximated per step by max landing words, [5] landing words,
	// [6] sum of max stored pixels, [7] stored pixels, [8] tasks, [9] lines.  For a candidate-at-a-time loop (each lane
	// either processes one candidate or, having none, takes its next DDA step): [10] sum over tasks of the MAX over lanes
	// of the lane's iterations (sum over its steps of max(1, candidates)), [11] sum over lanes of the same.
	template <bool kSse, bool kSse2> void warp_model(WalkIO wio, const osbn::Num<kSse> &num, const std::vector<ViewParams> &views,
	                                                 const std::vector<LineRec> &recs, const std::vector<uint32_t> &active,
	                                                 unsigned long long *out) {
		size_t a = 0;
		while (a < active.size()) {
			const int view = recs[active[a]].view;
			size_t b = a;
			while (b < active.size() && b - a < 32 && recs[active[b]].view == view)
				b++;
			const int n = (int)(b - a);
			std::vector<LineWalker<kSse, kSse2>> W(n);
			std::vector<std::vector<uint32_t>> filled(n, std::vector<uint32_t>(128 + kWalkExtraWords));
			std::vector<char> alive(n, 1);
			std::vector<unsigned long long> iters(n, 0);
			std::vector<int> nsteps(n, 0);
			int pred = 0;
			for (int i = 0; i < n; i++) {
				W[i].init(wio, active[a + i], (uint32_t)(a + i), filled[i].data(), 1);
				// prediction: distance at which the every-4th-step pitch cull fires (or the fog distance), times the cell
				// crossings per unit distance
				const float vmax = (float)wio.lineRes + 0.5f;
				float dEnd = 128.f;
				for (float dd = 0.5f; dd < 128.f; dd *= 1.09f) {
					const float id = 1.f / dd;
					if ((W[i].transform(num, id, 0, vmax) >= wio.lineRes - 1 && W[i].icz >= 0) || W[i].transform(num, id, 63, vmax) <= 0) {
						dEnd = dd;
						break;
					}
				}
				const float2 hd = wio.lineDirs[active[a + i]];
				if (getenv("EMUL_PRED_SCAN"))
					pred = std::max(pred, (int)(dEnd * (fabsf(hd.x) + fabsf(hd.y))));
				else
					pred = std::max(pred, (int)predict_steps(wio.lineTan[active[a + i]].x, wio.lineTan[active[a + i]].y, wio.lineRes,
					                                         views[view].origin[2], hd));
			}
			out[8]++;
			out[9] += n;
			double cost = 0;
			for (;;) {
				unsigned maxC = 0, sumC = 0, maxW = 0, sumW = 0, maxP = 0, sumP = 0, lanes = 0;
				for (int i = 0; i < n; i++) {
					if (!alive[i])
						continue;
					g_stat[0] = g_stat[1] = g_stat[2] = 0;
					alive[i] = W[i].step(wio, num, filled[i].data(), 1);
					lanes++;
					nsteps[i]++;
					maxC = std::max(maxC, g_stat[0]);
					sumC += g_stat[0];
					iters[i] += std::max(1u, g_stat[0]);
					maxW = std::max(maxW, g_stat[1]);
					sumW += g_stat[1];
					maxP = std::max(maxP, g_stat[2]);
					sumP += g_stat[2];
				}
				if (!lanes)
					break;
				out[0]++;
				out[1] += lanes;
				cost += 112 + 96 * maxC + 60 * maxW + 14 * maxP; // step + candidates + landing words + stored pixels

				out[2] += maxC;
				out[3] += sumC;
				out[4] += maxW;
				out[5] += sumW;
				out[6] += maxP;
				out[7] += sumP;
			}
			unsigned long long mx = 0;
			for (int i = 0; i < n; i++) {
				mx = std::max(mx, iters[i]);
				out[11] += iters[i];
			}
			out[10] += mx;
			g_taskCost.push_back((float)cost);
			g_taskView.push_back(view);
			g_taskSteps.push_back(*std::max_element(nsteps.begin(), nsteps.end()));
			g_taskPred.push_back(pred);
			a = b;
		}
	}
} // namespace

extern "C" {
// solid uint64[512][512] ([x][y]), color uint32[512][512][64]: the GameMap arrays; tiled here like osb200_map_upload
__attribute__((visibility("default"))) int emul_render(const uint64_t *solid, const uint32_t *color, const osb200_scene *scenes,
                                                       int nViews, int w, int h, const float fog[3], float fogDist, int numerics,
                                                       int threads, int under, int noMap, int noFog, uint32_t *fb, float *depth,
                                                       unsigned long long *stats3, int mapKey) {
	// mapKey: the tiled copy of the map is rebuilt whenever the key differs from the previous call's
	static std::vector<uint64_t> solidT;
	static std::vector<uint32_t> colorT;
	static int lastKey = -1;
	if (!noMap && (mapKey != lastKey || solidT.empty())) {
		solidT.resize(512 * 512);
		colorT.resize((size_t)512 * 512 * 64);
		for (uint32_t x = 0; x < 512; x++)
			for (uint32_t y = 0; y < 512; y++) {
				const uint32_t t = tiled_xy(x, y);
				solidT[t] = solid[x * 512 + y];
				std::memcpy(&colorT[(size_t)t * 64], &color[((size_t)x * 512 + y) * 64], 64 * 4);
			}
		lastKey = mapKey;
	}
	if (noMap && solidT.empty()) {
		solidT.assign(512 * 512, 0);
		colorT.assign((size_t)512 * 512 * 64, 0);
	}
	Stats st{};
	int rc;
#define RUN(SSE, SSE2) rc = run<SSE, SSE2>(solidT.data(), colorT.data(), scenes, nViews, w, h, fog, fogDist, threads, under, noMap, noFog, fb, depth, &st)
	if (numerics == 0)
		RUN(false, false);
	else if (numerics == 1)
		RUN(true, false);
	else
		RUN(true, true);
#undef RUN
	if (stats3) {
		stats3[0] = st.lines;
		stats3[1] = st.active;
		stats3[2] = st.steps;
	}
	return rc;
}
// costs of the tasks of the last emul_warp_model call (returns their number; cost / view may be null)
__attribute__((visibility("default"))) int emul_task_steps(int *steps, int *pred, int cap) {
	const int n = (int)g_taskSteps.size();
	for (int i = 0; i < n && i < cap; i++) {
		steps[i] = g_taskSteps[i];
		pred[i] = g_taskPred[i];
	}
	return n;
}
__attribute__((visibility("default"))) int emul_task_costs(float *cost, int *view, int cap) {
	const int n = (int)g_taskCost.size();
	for (int i = 0; i < n && i < cap; i++) {
		if (cost)
			cost[i] = g_taskCost[i];
		if (view)
			view[i] = g_taskView[i];
	}
	return n;
}
__attribute__((visibility("default"))) void emul_set_coop(int mode) { g_coop = mode; }
__attribute__((visibility("default"))) void emul_coop_stats(unsigned long long *out, int reset) {
	for (int i = 0; i < 4; i++) {
		out[i] = g_coopStats[i];
		if (reset)
			g_coopStats[i] = 0;
	}
}
// lock-step warp statistics of the walker for a batch of views (V-SSE2): out[10], see warp_model
__attribute__((visibility("default"))) int emul_warp_model(const uint64_t *solid, const uint32_t *color, const osb200_scene *scenes,
                                                           int nViews, int w, int h, unsigned long long *out) {
	std::vector<uint64_t> solidT(512 * 512);
	std::vector<uint32_t> colorT((size_t)512 * 512 * 64);
	for (uint32_t x = 0; x < 512; x++)
		for (uint32_t y = 0; y < 512; y++) {
			const uint32_t t = tiled_xy(x, y);
			solidT[t] = solid[x * 512 + y];
			std::memcpy(&colorT[(size_t)t * 64], &color[((size_t)x * 512 + y) * 64], 64 * 4);
		}
	const float fog[3] = {0.8f, 1.f, 1.f};
	for (int i = 0; i < 12; i++)
		out[i] = 0;
	g_taskCost.clear();
	g_taskView.clear();
	g_taskSteps.clear();
	g_taskPred.clear();
	return run<true, true>(solidT.data(), colorT.data(), scenes, nViews, w, h, fog, 128.f, 4, 1, 0, 0, nullptr, nullptr, nullptr, out);
}
}
