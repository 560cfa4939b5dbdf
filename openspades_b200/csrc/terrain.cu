// Software-renderer terrain pass + fog on the GPU, reproducing the reference's results:
//   SWRenderer::EndScene / BuildFrustrum   Sources/Draw/SWRenderer.cpp:207-227,709-745
//   SWMapRenderer::RenderInner             Sources/Draw/SWMapRenderer.cpp:1045-1174  (fill_view + k_view_prep)
//   SWMapRenderer::BuildLine               Sources/Draw/SWMapRenderer.cpp:219-671    (k_line_setup + k_walk_lines)
//   SWMapRenderer::RenderFinal             Sources/Draw/SWMapRenderer.cpp:744-1043   (k_gather_fog)
//   SWRenderer::ApplyFog                   Sources/Draw/SWRenderer.cpp:341-516       (fused into k_gather_fog)
//
// Design: the reference renders a view as ~0.6-5k vertical "yaw lines", each a 1-D strip of
// (colour, depth) built by a 2-D DDA through map columns, then gathers every screen pixel from (line, pitch index).
// The per-lane arithmetic lives in terrain_core.cuh (shared with the CPU emulation of tests/native); this file is
// the launch structure around it:
//   k_view_prep    one thread per (job, view): the reference's sequential float recurrences (line directions, fog / v1
//                  tables); a warp runs one recurrence for 32 views
//   k_line_setup   one thread per yaw line: pitch cull against the frustum (glibc-exact atanf / tanf), the line's pitch
//                  mapping, and per warp the ballot of the lines that survive
//   k_line_compact one CTA per view: order-preserving compaction of the survivors (a line's rank is its strip slot; culled
//                  lines own no strip memory), line records, and the walker's tasks: <= 32 adjacent lines of one view
//   k_walk_lines   persistent grid; a warp takes one task at a time from a batch-wide cursor, a lane walks one line (DDA +
//                  spans) in a loop of its own -- no launch tail per chunk, no lanes parked on culled lines
//   k_walk_coop    small launches (a single frame): one warp per line, see CoopLine in terrain_core.cuh
//   k_gather_fog   one warp per (view, 32-pixel column group, <= 8 block rows): RenderFinal + ApplyFog, frame written once
//   k_view_pipeline  optional (OSB200_PIPELINE): walk + gather of a view back to back in one persistent kernel; measured
//                  slower than the two kernels above
// All float arithmetic is separate IEEE mul/add (--fmad=false) in the reference's evaluation order.
#include "ctx.h"
#include "terrain_core.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace osb {

	// lower bound of lanes per warp for small k_walk_lines launches (OSB200_LINE_MIN_LANES, default 4; 32 = off).
	// One 1080p frame (3.8 k lines) with the round-1 kernel: 32 / 16 / 8 / 4 / 2 lanes -> 0.338 / 0.285 / 0.240 / 0.207 / 0.203 ms
	// OSB200_TASK_ORDER=1: walker tasks queued by estimated length, longest first (predict_steps, 16 queues) instead of
	// view order.  A list-scheduling model of the launch (tests/native: emul_task_costs) puts view order 10.6 % above the
	// ideal makespan and this order 2.9 % above it; on the B200 it is SLOWER (k_walk_lines 2.35 -> 2.59 ms per 4096
	// views, profiles/r02_ab_task_order.json): a view's neighbouring tasks no longer run side by side and stop sharing
	// their map columns and colours in L1 / L2, and the tail the model charges in full is mostly absorbed by the
	// remaining warps issuing faster.  Off by default; kept as the measured answer.
	static int task_order() {
		static const int v = [] {
			const char *e = getenv("OSB200_TASK_ORDER");
			return e ? atoi(e) : 0;
		}();
		return v;
	}
	// OSB200_PIPELINE=W: batches run k_view_pipeline (walk + gather of a view back to back in one persistent kernel, at most
	// W views between the first line walked and the last pixel gathered) instead of k_walk_lines + k_gather_fog.  0 = off.
	static int pipeline_window() {
		static const int v = [] {
			const char *e = getenv("OSB200_PIPELINE");
			return e ? std::max(0, atoi(e)) : 0;
		}();
		return v;
	}
	// OSB200_PIPE_SPLIT=k: of every 16 SMs, k walk and 16 - k gather (0 = every warp takes whatever is there)
	static int pipeline_split() {
		static const int v = [] {
			const char *e = getenv("OSB200_PIPE_SPLIT");
			return e ? std::min(15, std::max(0, atoi(e))) : 0;
		}();
		return v;
	}
	// OSB200_RCP_SHARED=0: k_walk_lines reads the rcpss table from global memory (A/B knob)
	static int rcp_shared() {
		static const int v = [] {
			const char *e = getenv("OSB200_RCP_SHARED");
			return e ? atoi(e) : 1;
		}();
		return v;
	}
	static int line_min_lanes() {
		static const int v = [] {
			const char *e = getenv("OSB200_LINE_MIN_LANES");
			const int n = e ? atoi(e) : 4;
			return n >= 32 ? 32 : (n >= 16 ? 16 : (n >= 8 ? 8 : (n >= 4 ? 4 : (n >= 2 ? 2 : 1))));
		}();
		return v;
	}
	// Per-chunk working set.  Two of them, each with its own stream, when a batch is cut into chunks (frames going
	// back to the host: the PCIe copy of chunk k overlaps the kernels of chunk k+1).
	struct ChunkBufs {
		float2 *d_lineDirs = nullptr;
		float2 *d_lineTan = nullptr;
		LineRec *d_lineRecs = nullptr;
		uint32_t *d_active = nullptr;   // strip slot -> line: the surviving lines of every view, compacted in yaw order
		uint32_t *d_liveBits = nullptr; // per view, per 32 lines: ballot of the lines that survived the pitch cull
		uint32_t *d_tasks = nullptr;    // walker tasks: (first slot << 5) | (lines - 1)
		uint32_t *d_counters = nullptr; // [0] number of tasks, [1] work cursor of k_walk_lines, [2..] fill + cursor per task queue
		uint32_t *d_bucketTasks = nullptr; // kTaskBuckets queues of bucketCap walker tasks (ordered mode)
		size_t bucketTaskCap = 0;
		uint32_t *d_tasksPV = nullptr; // k_view_pipeline: walker tasks in per-view regions of maxWordsPerView entries
		uint32_t *d_pipe = nullptr;    // ... and its counters: head | nTasks | taken | walked | gathered | ready queue (per view)
		size_t tasksPVCap = 0, pipeCap = 0;
		uint2 *d_strips = nullptr;
		uint32_t *d_lineMask = nullptr; // lineRes bits per strip slot: strip pixels the line wrote
		float *d_tabs = nullptr;
		size_t lineCap = 0, tanCap = 0, recCap = 0, activeCap = 0, liveCap = 0, taskCap = 0, stripCap = 0, maskCap = 0, tabCap = 0;
		cudaStream_t stream = nullptr;
		cudaEvent_t done = nullptr; // last kernel of the newest chunk that used this set
	};

	struct TerrainState {
		ViewParams *d_views = nullptr;
		ViewParams *h_views = nullptr; // pinned
		size_t viewCap = 0, dViewCap = 0, hTabCap = 0;
		ChunkBufs set[2];
		float *h_tabs = nullptr; // pinned (host-prep path)
		float2 *h_lineDirs = nullptr;
		size_t hLineCap = 0;
		uint16_t *d_atan = nullptr; // 4 x 5000: sm, lg, smN, lgN
		// persistent output buffers of the host-pointer entry point
		uint32_t *d_fb = nullptr;
		float *d_depth = nullptr;
		size_t outCap = 0;
		cudaEvent_t stageFree = nullptr;
		cudaStream_t copyStream = nullptr; // D2H of finished chunks, overlapped with the next chunk's kernels
		cudaEvent_t fork = nullptr, copyDone = nullptr;
		size_t walkSmem[3] = {0, 0, 0};
		int walkBlocksPerSm[3] = {0, 0, 0}; // resident CTAs of k_walk_lines per SM, per numerics variant (occupancy API)
	};

#ifndef OSB_WALK_THREADS
#define OSB_WALK_THREADS 128
#endif
#ifndef OSB_WALK_MINB
#define OSB_WALK_MINB 7
#endif
	constexpr int kWalkThreads = OSB_WALK_THREADS; // lanes (= lines in flight) per CTA of k_walk_lines
	constexpr int kSetupThreads = 128;
	constexpr int kMinTaskLines = 1; // smallest walker task (OSB200_LINE_MIN_LANES)
	constexpr int kCounterWords = 2 + 2 * kTaskBuckets;

	// Device-side recurrences for large batches: one warp per view, one lane per role.
	__global__ void k_view_prep(const ViewParams *__restrict__ views, int nViews, FrameGeom g, float2 *lineDirs, float *tabs) {
		// Every job is a sequential float recurrence (that is what makes the reference's tables what they are): one THREAD per
		// (job, view), the job in blockIdx.y, so that the 32 lanes of a warp run the same recurrence for 32 views.  (One warp
		// per view with a lane per job spent a whole warp instruction on every step of one lane's recurrence: 56 us.)
		//   job 0: line directions, 1: fog vx, 2 .. 2 + T - 1: fog vy band of emulated thread t, then T jobs: v1 columns
		const int vi = blockIdx.x * blockDim.x + threadIdx.x;
		if (vi >= nViews)
			return;
		const int job = blockIdx.y;
		const ViewParams &V = views[vi];
		float *vx = tabs + V.tabBase, *vy = vx + (g.fw >> 2), *v1 = vy + (g.fh >> 2);
		if (job == 0)
			prep_line_dirs(V, lineDirs + V.lineBase);
		else if (job == 1)
			prep_fog_vx(V, g, vx);
		else if (job < 2 + g.swThreads)
			prep_fog_vy(V, g, job - 2, vy);
		else
			prep_v1(V, g, job - 2 - g.swThreads, v1);
	}

	// One thread per yaw line (blockIdx.y = view): pitch cull against the frustum side planes.  Writes the line's
	// {pitchTanMin, pitchScale} and, per warp, the ballot of the lines that survive (liveBits[view * wordsPerView + w]).
	template <bool kSse>
	__global__ void __launch_bounds__(kSetupThreads)
	  k_line_setup(const ViewParams *__restrict__ views, const float2 *__restrict__ lineDirs, osbn::SseTables tabs, int lineRes,
	               int noMap, float2 *__restrict__ lineTan, uint32_t *__restrict__ liveBits, int wordsPerView) {
		const int vi = blockIdx.y;
		const ViewParams &V = views[vi];
		const int l = blockIdx.x * kSetupThreads + threadIdx.x;
		const bool valid = l < V.numLines;
		const uint32_t gl = V.lineBase + (uint32_t)(valid ? l : 0);
		LineSetupOut o;
		o.minTan = 0.f;
		o.pitchScale = 0.f;
		o.culled = true;
		if (valid && !noMap) {
			osbn::Num<kSse> num{tabs};
			o = line_setup<kSse>(V, lineDirs[gl], lineRes, num);
		}
		const unsigned live = __ballot_sync(0xffffffffu, valid && !o.culled);
		if ((threadIdx.x & 31) == 0)
			liveBits[(size_t)vi * wordsPerView + (l >> 5)] = live;
		if (valid)
			lineTan[gl] = make_float2(o.minTan, o.pitchScale);
	}

	// One CTA per view: order-preserving compaction of the view's surviving lines.  Their strip slots are
	// lineBase + rank (contiguous, in yaw order), the line records are written, and the view's lines are cut into
	// tasks of <= taskLines consecutive slots for the walker: a walker warp always gets ADJACENT lines of ONE view, which
	// walk almost the same cells (measured: with slots handed out in atomic order across views a warp mixed two or
	// three views and ran at 16.8 of 32 lanes).
	constexpr int kCompactThreads = 256;
	constexpr int kMaxLineWords = 65536 / 32; // numLines <= 65536 (SWMapRenderer.cpp:1116-1119)
	__global__ void __launch_bounds__(kCompactThreads)
	  k_line_compact(const ViewParams *__restrict__ views, const float2 *__restrict__ lineTan, const uint32_t *__restrict__ liveBits,
	                 int wordsPerView, int taskLines, LineRec *__restrict__ lineRecs, uint32_t *__restrict__ active,
	                 uint32_t *__restrict__ tasks, uint32_t *__restrict__ counters, const float2 *__restrict__ lineDirs, int lineRes,
	                 uint32_t *__restrict__ bucketTasks, uint32_t bucketCap, uint32_t *__restrict__ tasksPV,
	                 uint32_t *__restrict__ nTasksPV, int maxTasksPV) {
		__shared__ uint32_t s_prefix[kMaxLineWords]; // exclusive prefix of the survivor counts per 32-line word
		__shared__ uint32_t s_warpSum[kCompactThreads / 32];
		__shared__ uint32_t s_base, s_taskBase;
		// ordered mode (bucketTasks != nullptr, tasks of 32 lines): longest estimated line per task, bucket fill counts
		__shared__ int s_taskSteps[kMaxLineWords];
		__shared__ uint32_t s_hist[kTaskBuckets], s_bucketBase[kTaskBuckets];
		const bool ordered = bucketTasks != nullptr;
		const int vi = blockIdx.x;
		const ViewParams &V = views[vi];
		const int nWords = (V.numLines + 31) >> 5;
		const uint32_t *bits = liveBits + (size_t)vi * wordsPerView;
		const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
		if (threadIdx.x == 0)
			s_base = 0;
		__syncthreads();
		for (int c = 0; c < nWords; c += kCompactThreads) { // block scan, kCompactThreads words at a time
			const int j = c + threadIdx.x;
			const uint32_t cnt = j < nWords ? (uint32_t)__popc(bits[j]) : 0u;
			uint32_t incl = cnt;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
				if (lane >= d)
					incl += up;
			}
			if (lane == 31)
				s_warpSum[wid] = incl;
			__syncthreads();
			uint32_t before = s_base;
			for (int w = 0; w < wid; w++)
				before += s_warpSum[w];
			if (j < nWords)
				s_prefix[j] = before + incl - cnt;
			__syncthreads();
			if (threadIdx.x == kCompactThreads - 1)
				s_base = before + incl;
			__syncthreads();
		}
		const uint32_t nLive = s_base;
		if (ordered) {
			for (int i = threadIdx.x; i < nWords; i += kCompactThreads)
				s_taskSteps[i] = 0;
			if (threadIdx.x < kTaskBuckets)
				s_hist[threadIdx.x] = 0;
			__syncthreads();
		}
		for (int j = wid; j < nWords; j += kCompactThreads / 32) { // one warp per word: lane = line
			const uint32_t b = bits[j];
			const int l = (j << 5) + lane;
			if (l >= V.numLines)
				continue;
			const uint32_t gl = V.lineBase + (uint32_t)l;
			const float2 tn = lineTan[gl];
			LineSetupOut o;
			o.minTan = tn.x;
			o.pitchScale = tn.y;
			o.culled = !((b >> lane) & 1u);
			int slot = -1;
			if (!o.culled) {
				const uint32_t rank = s_prefix[j] + (uint32_t)__popc(b & ((1u << lane) - 1u));
				slot = (int)(V.lineBase + rank);
				active[slot] = gl;
				if (ordered)
					atomicMax(&s_taskSteps[rank >> 5], (int)predict_steps(tn.x, tn.y, lineRes, V.origin[2], lineDirs[gl]));
			}
			// (a culled line never reaches the walker: its record carries the empty hull from the start)
			lineRecs[gl] = make_line_rec(o, slot, (OSB_STRIP_HULL && slot < 0) ? kEmptyHull : vi, lineRes);
		}
		const uint32_t nTasks = (nLive + (uint32_t)taskLines - 1) / (uint32_t)taskLines;
		if (threadIdx.x == 0)
			s_taskBase = nTasks ? atomicAdd(counters, nTasks) : 0u;
		__syncthreads();
		if (tasksPV) { // k_view_pipeline: the view's tasks in a region of their own, taken view by view
			for (uint32_t i = threadIdx.x; i < nTasks; i += kCompactThreads) {
				const uint32_t first = i * (uint32_t)taskLines;
				const uint32_t n = min((uint32_t)taskLines, nLive - first);
				tasksPV[(size_t)vi * maxTasksPV + i] = ((V.lineBase + first) << 5) | (n - 1u);
			}
			if (threadIdx.x == 0)
				nTasksPV[vi] = nTasks;
			return;
		}
		if (!ordered) {
			for (uint32_t i = threadIdx.x; i < nTasks; i += kCompactThreads) {
				const uint32_t first = i * (uint32_t)taskLines;
				const uint32_t n = min((uint32_t)taskLines, nLive - first);
				tasks[s_taskBase + i] = ((V.lineBase + first) << 5) | (n - 1u); // slot of the first line, lines - 1
			}
			return;
		}
		// ordered: the view's tasks go into the batch-wide queues by estimated length (one global atomic per non-empty
		// bucket per view)
		for (uint32_t i0 = 0; i0 < nTasks; i0 += kCompactThreads) { // nTasks <= 2048: a few rounds at most
			const uint32_t i = i0 + threadIdx.x;
			int bkt = 0;
			uint32_t pos = 0;
			if (i < nTasks) {
				bkt = task_bucket((float)s_taskSteps[i]);
				pos = atomicAdd(&s_hist[bkt], 1u);
			}
			__syncthreads();
			if (threadIdx.x < kTaskBuckets) {
				const uint32_t cnt = s_hist[threadIdx.x];
				s_bucketBase[threadIdx.x] = cnt ? atomicAdd(counters + 2 + threadIdx.x, cnt) : 0u;
			}
			__syncthreads();
			if (i < nTasks) {
				const uint32_t first = i * 32u;
				const uint32_t n = min(32u, nLive - first);
				bucketTasks[(size_t)bkt * bucketCap + s_bucketBase[bkt] + pos] = ((V.lineBase + first) << 5) | (n - 1u);
			}
			__syncthreads();
			if (threadIdx.x < kTaskBuckets)
				s_hist[threadIdx.x] = 0;
			__syncthreads();
		}
	}

	// Persistent line walker.  A warp takes one task (<= 32 adjacent surviving lines of one view) at a time from the
	// chunk-wide cursor; lane = one yaw line.
	// kRcpS: the rcpss table (<= kRcpSharedEntries entries) is copied to shared memory once per CTA -- the walker looks it
	// up once per DDA step and twice per span candidate, each time with an immediately dependent use
	constexpr int kRcpSharedBits = 11, kRcpSharedEntries = 1 << kRcpSharedBits;
	template <bool kSse, bool kSse2Shade, bool kRcpS = false>
	__global__ void __launch_bounds__(kWalkThreads, OSB_WALK_MINB)
	  k_walk_lines(WalkIO io, osbn::SseTables tabs, const uint32_t *__restrict__ active, const uint32_t *__restrict__ tasks,
	               uint32_t *__restrict__ counters, uint32_t bucketCap) {
		extern __shared__ uint32_t s_filled[]; // (lineRes/32 + kWalkExtraWords) words x kWalkThreads, word-major: bank = lane
		__shared__ uint32_t s_rcp[kRcpS ? kRcpSharedEntries : 1];
		uint32_t *filled = s_filled + threadIdx.x;
		const int lane = threadIdx.x & 31;
		if (kRcpS) {
			for (int i = threadIdx.x; i < (1 << tabs.rcpBits); i += kWalkThreads)
				s_rcp[i] = __ldg(tabs.rcp + i);
			__syncthreads();
			tabs.rcp = s_rcp;
		}
		const osbn::Num<kSse> num{tabs};
		const uint32_t nTasks = counters[0]; // written by k_line_compact, earlier on this stream
		// bucketCap != 0: `tasks` holds kTaskBuckets queues of bucketCap entries, filled to counters[2 + b], taken from
		// counters[2 + kTaskBuckets + b]; the warp drains the queue of the longest lines first (longest-processing-time
		// order: the launch ends with short tasks, not with one warp in the middle of a 180-step line)
		int bucket = kTaskBuckets - 1;
		LineWalker<kSse, kSse2Shade> W;
		for (;;) {
			uint32_t t = 0;
			if (bucketCap) {
				if (lane == 0) {
					for (;;) {
						t = atomicAdd(counters + 2 + kTaskBuckets + bucket, 1u);
						if (t < counters[2 + bucket])
							break;
						if (--bucket < 0)
							break;
					}
				}
				bucket = __shfl_sync(0xffffffffu, bucket, 0);
				if (bucket < 0)
					break;
				t = __shfl_sync(0xffffffffu, t, 0) + (uint32_t)bucket * bucketCap;
			} else {
				if (lane == 0)
					t = atomicAdd(counters + 1, 1u);
				t = __shfl_sync(0xffffffffu, t, 0);
				if (t >= nTasks)
					break;
			}
			const uint32_t task = __ldg(tasks + t);
			const uint32_t slot = (task >> 5) + (uint32_t)lane;
			bool alive = (uint32_t)lane <= (task & 31u);
			if (alive) {
				W.init(io, __ldg(active + slot), slot, filled, kWalkThreads);
				// every lane walks its own line to the end; the warp meets again behind the loop (a warp-wide vote per step
				// cost a dozen instructions of votes and register shuffles per DDA step)
				while (W.step(io, num, filled, kWalkThreads)) {
				}
			}
			__syncwarp();
		}
	}

	// Cooperative walker for small launches: one warp (= one CTA) per yaw line, see CoopLine in terrain_core.cuh.
	// Tasks are single lines (k_line_compact with taskLines = 1).
	template <bool kSse, bool kSse2Shade>
	__global__ void __launch_bounds__(32)
	  k_walk_coop(WalkIO io, osbn::SseTables tabs, const uint32_t *__restrict__ active, const uint32_t *__restrict__ tasks,
	              const uint32_t *__restrict__ counters) {
		extern __shared__ uint32_t s_coop[]; // CoopShared, then the strip's pixel keys
		CoopShared &sh = *reinterpret_cast<CoopShared *>(s_coop);
		uint32_t *key = s_coop + sizeof(CoopShared) / sizeof(uint32_t);
		const int lane = threadIdx.x;
		const osbn::Num<kSse> num{tabs};
		const uint32_t nTasks = counters[0];
		for (uint32_t t = blockIdx.x; t < nTasks; t += gridDim.x) {
			const uint32_t slot = __ldg(tasks + t) >> 5;
			const uint32_t gl = __ldg(active + slot);
			CoopLine<kSse, kSse2Shade> C;
			C.begin(io, gl, slot, sh, key, lane);
			__syncwarp();
			if (lane == 0)
				C.path(sh);
			__syncwarp();
			const int nSteps = sh.nSteps;
			bool ended = false;
			for (int s0 = 0; s0 < nSteps && !ended; s0 += 32) {
				const int s = s0 + lane;
				const unsigned cull = __ballot_sync(0xffffffffu, C.culls_after(io, num, sh, s));
				const int cullAt = cull ? s0 + __ffs((int)cull) - 1 : 0x7fffffff; // the step after which the line ends
				if (s < nSteps && s <= cullAt)
					C.emit_step(io, num, sh, key, s);
				ended = cull != 0u;
			}
			__syncwarp();
			if (sh.nSpans > kCoopMaxSpans) { // too many spans to record: this line goes through the sequential walker
				if (lane == 0) {
					LineWalker<kSse, kSse2Shade> W;
					W.init(io, gl, slot, key, 1);
					while (W.step(io, num, key, 1)) {
					}
				}
				__syncwarp();
				continue;
			}
			C.resolve_long(sh, key, lane);
			__syncwarp();
			int first, last;
			C.scan_hull(io, key, lane, first, last);
			first = __reduce_min_sync(0xffffffffu, first);
			last = __reduce_max_sync(0xffffffffu, last);
			C.store(io, sh, key, lane, first, last);
			__syncwarp();
		}
	}

	// RenderFinal + ApplyFog, see GatherLane in terrain_core.cuh.
	// OSB_GATHER_TMA=1 (experiment): the CTA stages its view's line-record table (numLines x 16 B)
	// in shared memory with ONE bulk asynchronous copy (cp.async.bulk, the 1-D TMA path, completion on an mbarrier) while
	// the warps run their fog / corner set-up, and the per-pixel record loads become shared-memory loads.
#ifndef OSB_GATHER_TMA
#define OSB_GATHER_TMA 0
#endif
#if OSB_GATHER_TMA
	constexpr int kStagedRecs = 1536; // line records a CTA can stage (24 KB): covers 256x256 views (<= 1192 lines)
#endif
#ifndef OSB_GATHER_MINB
#define OSB_GATHER_MINB 9 // <= 56 registers, no spills: 1.15 ms per 4096 views; 8 CTAs (62 registers) 1.17, 10 CTAs (48) 1.25
#endif
	template <bool kSse, bool kSse2Fog>
	__global__ void __launch_bounds__(kGatherWarps * 32, OSB_GATHER_MINB) k_gather_fog(GatherIO io, osbn::SseTables tabs) {
		__shared__ GatherShared s_g[kGatherWarps];
		const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
		const uint32_t task = blockIdx.x * kGatherWarps + wid;
		const osbn::Num<kSse> num{tabs};
		GatherShared &sm = s_g[wid];
		GatherLane<kSse, kSse2Fog> L;
		const bool mine = L.init(io, task, lane); // false: the task lies outside the launch (whole warp)
#if OSB_GATHER_TMA
		__shared__ alignas(128) LineRec s_recs[kStagedRecs];
		__shared__ alignas(8) unsigned long long s_bar;
		// the view of the CTA's first task; its table is staged when it fits (warps of another view read global memory)
		const uint32_t perView = (uint32_t)((((io.g.fw >> 3) + 3) >> 2) * (((io.g.fh >> 3) + io.segRows - 1) / io.segRows));
		const uint32_t vi0 = (blockIdx.x * kGatherWarps) / perView;
		const ViewParams *V0 = vi0 < (uint32_t)io.nViews ? io.views + vi0 : nullptr;
		const bool stage = V0 && V0->numLines <= kStagedRecs;
		const uint32_t barAddr = (uint32_t)__cvta_generic_to_shared(&s_bar);
		if (threadIdx.x == 0 && stage) {
			asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barAddr));
			asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		}
		__syncthreads();
		if (threadIdx.x == 0 && stage) {
			const uint32_t bytes = (uint32_t)V0->numLines * (uint32_t)sizeof(LineRec);
			const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_recs);
			const LineRec *src = io.lineRecs + V0->lineBase;
			asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barAddr), "r"(bytes) : "memory");
			asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
			             "l"(src), "r"(bytes), "r"(barAddr)
			             : "memory");
		}
#endif
		if (!mine)
			return;
		L.fog_phase(io, num, sm, lane);
		__syncwarp();
#if OSB_GATHER_TMA
		if (stage && L.V == V0) { // wait for the bulk copy (phase 0 of the barrier), then read records from shared memory
			uint32_t done = 0;
			while (!done)
				asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}"
				             : "=r"(done)
				             : "r"(barAddr)
				             : "memory");
			L.stagedRecs = s_recs;
		}
#endif
		for (int by = L.by0; by < L.by1; by += 2) {
			L.corner_phase(io, num, sm, lane, by);
			__syncwarp();
			OSB_PRAGMA(unroll OSB_GATHER_ROW_UNROLL)
			for (int r = 0; r < 2; r++)
				L.pixel_phase(io, num, sm, r, by);
			__syncwarp();
		}
	}

	// ------------------------------------------------------------------------------------------------------------------
	// View-granular pipeline: ONE persistent kernel walks the lines of a view and gathers + fogs that view as soon as its
	// last line is done, so that the view's strips are read back while they are still in L2 instead of making the round
	// trip through HBM that two batch-wide kernels force (walker writes 1.8 GB of strips, gather reads 1.6 GB back).
	// A warp is a walker or a gatherer task by task:
	//   gather first   a ticket of the oldest walked view (ready queue, in the order the views' walks completed; the
	//                  tickets on offer are counted by a semaphore);
	//   else walk      the next task of the current view -- views are admitted in index order, and view v only once view
	//                  v - window has been gathered completely: the strips in flight stay within `window` views;
	//   else           back off and look again.
	// Hand-over of a view from its walker warps to its gatherer warps: every lane fences its strip / hull stores, lane 0
	// counts the task done with an atomic; the warp that counts the last one publishes the view in the ready queue; the
	// taker reads the queue entry, fences, and reads strips and line records with ld.global.cg (L1 is not coherent).
	// ------------------------------------------------------------------------------------------------------------------
	constexpr int kPipeCur = 0, kPipeReady = 32, kPipeGather = 64, kPipeAvail = 16; // hot counters on separate lines / sectors
	constexpr int kPipeHead = 128;                                 // words before the per-view arrays
	__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t *p) {
		uint32_t v;
		asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
		return v;
	}
	template <bool kSse, bool kSse2>
	__global__ void __launch_bounds__(kWalkThreads, OSB_WALK_MINB)
	  k_view_pipeline(WalkIO wio, GatherIO gio, osbn::SseTables tabs, const uint32_t *__restrict__ active,
	                  const uint32_t *__restrict__ tasksPV, int maxTasksPV, uint32_t *__restrict__ pipe, uint32_t window,
	                  uint32_t perView, uint32_t split) {
		extern __shared__ uint32_t s_filled[]; // the walker's fill masks, as in k_walk_lines
		__shared__ GatherShared s_g[kWalkThreads / 32];
		uint32_t *filled = s_filled + threadIdx.x;
		const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
		const osbn::Num<kSse> num{tabs};
		const uint32_t nViews = (uint32_t)gio.nViews, totalG = nViews * perView;
		const uint32_t *nTasksPV = pipe + kPipeHead;
		uint32_t *viewTaken = pipe + kPipeHead + nViews, *walkDone = viewTaken + nViews, *gatherDone = walkDone + nViews,
		         *readyQ = gatherDone + nViews;
		// split != 0: roles per SM instead of per warp -- SMs with smid % 16 < split walk (and gather only once every walk task
		// is handed out), the others gather (and walk only until the first view is ready): an SM then runs ONE of the two
		// instruction streams at a time (together they do not fit the instruction cache)
		uint32_t smid = 0;
		asm("mov.u32 %0, %%smid;" : "=r"(smid));
		const bool mayGatherFirst = split == 0 || (smid & 15u) >= split;
		const bool walkerSm = split == 0 || (smid & 15u) < split;
		auto publish = [&](uint32_t v) { // lane 0, after its fence
			const uint32_t idx = atomicAdd(pipe + kPipeReady, 1u);
			asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(readyQ + idx), "r"(v) : "memory");
			__threadfence();
			atomicAdd(reinterpret_cast<int *>(pipe + kPipeAvail), (int)perView); // perView more gather tickets
		};
#ifdef OSB_PIPE_STATS
		unsigned long long tDisp = 0, tWalk = 0, tGather = 0, nPoll = 0, nSleep = 0, nCasFail = 0;
#define PIPE_T0 const long long t0_ = clock64()
#define PIPE_ACC(x) x += (unsigned long long)(clock64() - t0_)
#else
#define PIPE_T0
#define PIPE_ACC(x)
#endif
		for (;;) {
			uint32_t kind = 0, a = 0, b = 0; // 1 walk (view a, task b), 2 gather (ticket a), 3 all work handed out
			if (lane == 0) {
				PIPE_T0;
				uint32_t backoff = 64;
				for (;;) {
					if (ld_relaxed(pipe + kPipeGather) >= totalG) {
						kind = 3;
						break;
					}
					// gather tickets are a semaphore (fetch-and-add only: a compare-and-swap ticket collapses to one success
					// per L2 round trip once a few thousand warps contend -- measured, 90x slower)
					const uint32_t cur = ld_relaxed(pipe + kPipeCur);
					if ((mayGatherFirst || cur >= nViews) && (int)ld_relaxed(pipe + kPipeAvail) > 0) {
						if (atomicSub(reinterpret_cast<int *>(pipe + kPipeAvail), 1) > 0) {
							kind = 2;
							a = atomicAdd(pipe + kPipeGather, 1u);
							break;
						}
						atomicAdd(reinterpret_cast<int *>(pipe + kPipeAvail), 1); // none left after all: give it back
					}
					const uint32_t v = cur;
					if ((walkerSm || ld_relaxed(pipe + kPipeReady) == 0u) && v < nViews && (v < window || ld_relaxed(gatherDone + (v - window)) >= perView)) {
						const uint32_t nt = nTasksPV[v];
						if (ld_relaxed(viewTaken + v) < nt) { // (looked at first: a herd of idle warps must not pile atomics on it)
							const uint32_t t = atomicAdd(viewTaken + v, 1u);
							if (t < nt) {
								kind = 1;
								a = v;
								b = t;
								break;
							}
						}
						// the view's tasks are all handed out: move on (a view without surviving lines is ready at once)
						if (atomicCAS(pipe + kPipeCur, v, v + 1u) == v && nt == 0u)
							publish(v);
						continue;
					}
#ifdef OSB_PIPE_STATS
					nSleep++;
#endif
					__nanosleep(backoff);
					if (backoff < 2048)
						backoff <<= 1;
				}
				PIPE_ACC(tDisp);
			}
			kind = __shfl_sync(0xffffffffu, kind, 0);
			a = __shfl_sync(0xffffffffu, a, 0);
			b = __shfl_sync(0xffffffffu, b, 0);
			if (kind == 3)
				break;
			if (kind == 1) {
				PIPE_T0;
				const uint32_t task = __ldg(tasksPV + (size_t)a * maxTasksPV + b);
				const uint32_t slot = (task >> 5) + (uint32_t)lane;
				bool alive = (uint32_t)lane <= (task & 31u);
				LineWalker<kSse, kSse2> W;
				if (alive)
					W.init(wio, __ldg(active + slot), slot, filled, kWalkThreads);
				while (__any_sync(0xffffffffu, alive)) {
					if (alive)
						alive = W.step(wio, num, filled, kWalkThreads);
				}
				__threadfence(); // this lane's strip pixels and hull, before the task is counted
				__syncwarp();
				if (lane == 0) {
					__threadfence();
					if (atomicAdd(walkDone + a, 1u) + 1u == nTasksPV[a]) {
						__threadfence();
						publish(a);
					}
					PIPE_ACC(tWalk);
				}
			} else {
				PIPE_T0;
				const uint32_t ridx = a / perView, sub = a - ridx * perView;
				uint32_t v = 0;
				if (lane == 0) {
					do
						v = ld_relaxed(readyQ + ridx); // reserved before it was written: a few cycles at most
					while (v == 0xffffffffu);
					__threadfence();
				}
				v = __shfl_sync(0xffffffffu, v, 0);
				GatherShared &sm = s_g[wid];
				GatherLane<kSse, kSse2, true> L;
				L.init(gio, v * perView + sub, lane);
				L.fog_phase(gio, num, sm, lane);
				__syncwarp();
				for (int by = L.by0; by < L.by1; by += 2) {
					L.corner_phase(gio, num, sm, lane, by);
					__syncwarp();
					OSB_PRAGMA(unroll 1)
					for (int r = 0; r < 2; r++)
						L.pixel_phase(gio, num, sm, r, by);
					__syncwarp();
				}
				if (lane == 0) {
					atomicAdd(gatherDone + v, 1u);
					PIPE_ACC(tGather);
				}
			}
		}
#ifdef OSB_PIPE_STATS
		if (lane == 0) {
			unsigned long long *st = reinterpret_cast<unsigned long long *>(pipe + 96);
			atomicAdd(st + 0, tDisp);
			atomicAdd(st + 1, tWalk);
			atomicAdd(st + 2, tGather);
			atomicAdd(st + 3, nSleep);
		}
#endif
	}

	static int terrain_init(osb200_ctx *c) {
		if (c->terrain)
			return 0;
		auto *T = new TerrainState();
		c->terrain = T;
		// AtanTable (SWMapRenderer.cpp:673-696), built with the host libm exactly like the reference
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
		OSB_CUDA(cudaMalloc(&T->d_atan, tab.size() * 2));
		OSB_CUDA(cudaMemcpy(T->d_atan, tab.data(), tab.size() * 2, cudaMemcpyHostToDevice));
		OSB_CUDA(cudaEventCreateWithFlags(&T->stageFree, cudaEventDisableTiming));
		OSB_CUDA(cudaStreamCreateWithFlags(&T->copyStream, cudaStreamNonBlocking));
		OSB_CUDA(cudaEventCreateWithFlags(&T->fork, cudaEventDisableTiming));
		OSB_CUDA(cudaEventCreateWithFlags(&T->copyDone, cudaEventDisableTiming));
		for (ChunkBufs &b : T->set) {
			OSB_CUDA(cudaStreamCreateWithFlags(&b.stream, cudaStreamNonBlocking));
			OSB_CUDA(cudaEventCreateWithFlags(&b.done, cudaEventDisableTiming));
		}
		return 0;
	}

	void terrain_destroy(osb200_ctx *c) {
		TerrainState *T = c->terrain;
		if (!T)
			return;
		cudaFree(T->d_views);
		cudaFreeHost(T->h_views);
		for (ChunkBufs &b : T->set) {
			cudaFree(b.d_lineDirs);
			cudaFree(b.d_lineTan);
			cudaFree(b.d_lineRecs);
			cudaFree(b.d_active);
			cudaFree(b.d_liveBits);
			cudaFree(b.d_tasks);
			cudaFree(b.d_counters);
			cudaFree(b.d_bucketTasks);
			cudaFree(b.d_tasksPV);
			cudaFree(b.d_pipe);
			cudaFree(b.d_strips);
			cudaFree(b.d_lineMask);
			cudaFree(b.d_tabs);
			if (b.done)
				cudaEventDestroy(b.done);
			if (b.stream)
				cudaStreamDestroy(b.stream);
		}
		cudaFreeHost(T->h_tabs);
		cudaFreeHost(T->h_lineDirs);
		cudaFree(T->d_atan);
		cudaFree(T->d_fb);
		cudaFree(T->d_depth);
		if (T->stageFree)
			cudaEventDestroy(T->stageFree);
		if (T->fork)
			cudaEventDestroy(T->fork);
		if (T->copyDone)
			cudaEventDestroy(T->copyDone);
		if (T->copyStream)
			cudaStreamDestroy(T->copyStream);
		delete T;
		c->terrain = nullptr;
	}

	template <class T> static int grow(T *&p, size_t &cap, size_t want, bool pinned = false) {
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
		size_t n = want + (want >> 3);
		if (pinned)
			OSB_CUDA(cudaMallocHost((void **)&p, n * sizeof(T)));
		else
			OSB_CUDA(cudaMalloc((void **)&p, n * sizeof(T)));
		cap = n;
		return 0;
	}

	static size_t env_mb(const char *name, long dflt) {
		const char *e = getenv(name);
		long mb = e ? atol(e) : 0;
		return (size_t)(mb > 0 ? mb : dflt) << 20;
	}
	// Strip bytes per chunk, sized for the worst case (every yaw line of the chunk survives the pitch cull; on the
	// headline batch about 40 % do, and only those pages are ever touched).  A 4096-view batch of 256x256 is 6.7 GB of
	// strip address space: one chunk on a 180 GB part.
	static size_t strip_budget_bytes() {
		static size_t v = env_mb("OSB200_STRIP_MB", 12288);
		return v;
	}
	// experiment knobs: OSB200_MIN_CHUNKS cuts a large device-output batch into at least that many chunks (alternating on
	// two streams: chunk k's gather runs beside chunk k+1's walker); OSB200_WALK_BLOCKS caps the walker's resident CTAs per
	// SM so that gather CTAs of the other stream find registers on the same SM.
	static int min_chunks() {
		static const int v = [] {
			const char *e = getenv("OSB200_MIN_CHUNKS");
			const int n = e ? atoi(e) : 1;
			return n < 1 ? 1 : n;
		}();
		return v;
	}
	static int walk_blocks_cap() {
		static const int v = [] {
			const char *e = getenv("OSB200_WALK_BLOCKS");
			const int n = e ? atoi(e) : 0;
			return n < 0 ? 0 : n;
		}();
		return v;
	}
	// OSB200_COOP: 1 = always the cooperative walker, 0 = never, unset = by launch size
	static int coop_mode() {
		static const int v = [] {
			const char *e = getenv("OSB200_COOP");
			return e ? (atoi(e) != 0 ? 1 : 0) : -1;
		}();
		return v;
	}
	static bool dual_stream() {
		static int v = -1;
		if (v < 0) {
			const char *e = getenv("OSB200_DUAL_STREAM");
			v = e ? atoi(e) != 0 : 1;
		}
		return v != 0;
	}

	template <class K> static int resident_blocks(K kernel, int threads, size_t smem) {
		int nb = 0;
		if (smem > 48 * 1024) // strips of 2048+ pixels (frames 2048+ rows high)
			cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, threads, smem) != cudaSuccess || nb < 1)
			nb = 1;
		return nb;
	}

	// host_fb / host_depth (nullable): when given, each chunk's finished frames are copied to the host on a second
	// stream while the next chunk renders (PCIe overlaps the kernels); the caller synchronises copyDone.
	int render_impl(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
	                float fogDist, uint32_t *d_fb, float *d_depth, uint32_t *host_fb, float *host_depth, bool deferFog) {
		if (!c) {
			set_error("osb200_render: no context");
			return -1;
		}
		if (nViews == 0)
			return 0;
		if (!scenes || !fogColor || !d_fb || !d_depth || nViews < 0) {
			set_error("osb200_render: null argument");
			return -1;
		}
		if (w <= 0 || h <= 0 || (w & 7) || (h & 7)) {
			set_error("Framebuffer size is not multiple of 8."); // SWRenderer.cpp:105-107
			return -1;
		}
		if (h >= 8192 || w > 32768) {
			set_error("osb200_render: framebuffer too large (h < 8192, w <= 32768)");
			return -1;
		}
		OSB_CUDA(cudaSetDevice(c->device));
		if (terrain_init(c))
			return -2;
		TerrainState *T = c->terrain;
		const int under = c->opts.undersampling;
		FrameGeom g;
		g.fw = w;
		g.fh = h;
		g.under = under;
		g.swThreads = c->opts.sw_threads;
		g.sse2Fog = g.sse2Shade = c->opts.numerics == OSB200_NUM_SSE2;
		g.noMap = c->mapLoaded ? 0 : 1;
		g.noFog = deferFog ? 1 : 0;
		auto toFixed8 = [](float v) { // SWUtils.h:68-71
			int i = static_cast<int>(v * 255.f + .5f);
			return (uint32_t)std::max(std::min(i, 255), 0);
		};
		g.fogR = toFixed8(fogColor[0]);
		g.fogG = toFixed8(fogColor[1]);
		g.fogB = toFixed8(fogColor[2]);

		// the pinned staging buffers may still be in flight from the previous call
		OSB_CUDA(cudaEventSynchronize(T->stageFree));
		if (grow(T->h_views, T->viewCap, (size_t)nViews, true))
			return -2;
		if (grow(T->d_views, T->dViewCap, (size_t)nViews))
			return -2;
		const uint32_t tabF = tab_floats(w, h);
		int lineRes = 0;
		for (int i = 0; i < nViews; i++) {
			fill_view(scenes[i], w, h, under, fogDist, T->h_views[i]);
			lineRes = T->h_views[i].lineRes;
		}
		const int words = lineRes >> 5;
		// ---- chunking: the strips of one chunk must fit the budget; chunks are made equal ----
		size_t budgetLines = std::max<size_t>(strip_budget_bytes() / ((size_t)lineRes * 8), 65536);
		{
			size_t total = 0;
			for (int i = 0; i < nViews; i++)
				total += (size_t)T->h_views[i].numLines;
			size_t want = (total + budgetLines - 1) / budgetLines;
			// frames going back to the host: the PCIe copy (4-8 B/pixel) is the long pole, so start it early --
			// more, smaller chunks cost kernel time that hides under the copy of the previous chunk
			if ((host_fb || host_depth) && nViews >= 1024 && want < 8)
				want = 8;
			if (nViews >= 512 && want < (size_t)min_chunks())
				want = (size_t)min_chunks();
			budgetLines = std::max<size_t>((total + want - 1) / want + 65536 / 32, 65536);
		}
		std::vector<int> chunkStart;
		size_t maxLines = 0, maxViews = 0;
		{
			size_t lines = 0;
			int start = 0;
			chunkStart.push_back(0);
			for (int i = 0; i < nViews; i++) {
				ViewParams &v = T->h_views[i];
				if (lines + (size_t)v.numLines > budgetLines && i > start) {
					chunkStart.push_back(i);
					start = i;
					lines = 0;
				}
				v.lineBase = (uint32_t)lines;
				v.tabBase = (uint32_t)(i - start) * tabF;
				lines += (size_t)v.numLines;
				maxLines = std::max(maxLines, lines);
				maxViews = std::max(maxViews, (size_t)(i - start + 1));
			}
			chunkStart.push_back(nViews);
		}
		if (maxLines * (size_t)lineRes >= ((size_t)1 << 32)) { // strip pixels are indexed with 32 bits inside a chunk
			set_error("osb200_render: OSB200_STRIP_MB too large for %d-pixel strips (a chunk must stay below 2^32 strip pixels)", lineRes);
			return -1;
		}
		int maxViewLinesAll = 0;
		for (int i = 0; i < nViews; i++)
			maxViewLinesAll = std::max(maxViewLinesAll, T->h_views[i].numLines);
		// words of the per-view survivor ballots: the k_line_setup grid covers whole CTAs of lines
		const int maxWordsPerView = (maxViewLinesAll + kSetupThreads - 1) / kSetupThreads * (kSetupThreads / 32);
		const bool hostPrep = nViews <= 8;
		const size_t nChunks = chunkStart.size() - 1;
		// one chunk (or the host-prep path with its shared pinned staging): everything on the context stream
		// (per-kernel event timing wants each kernel alone on the GPU: with timing on, chunks run back to back)
		const bool dual = dual_stream() && nChunks > 1 && !hostPrep && !c->timing;
		for (int k = 0; k < (dual ? 2 : 1); k++) {
			ChunkBufs &B = T->set[k];
			if (grow(B.d_lineDirs, B.lineCap, maxLines) || grow(B.d_lineTan, B.tanCap, maxLines) ||
			    grow(B.d_lineRecs, B.recCap, maxLines) || grow(B.d_active, B.activeCap, maxLines) ||
			    grow(B.d_liveBits, B.liveCap, maxViews * (size_t)maxWordsPerView) ||
			    grow(B.d_tasks, B.taskCap, maxLines / kMinTaskLines + maxViews) ||
			    grow(B.d_bucketTasks, B.bucketTaskCap, (size_t)kTaskBuckets * (maxLines / 32 + maxViews)) ||
			    (pipeline_window() &&
			     (grow(B.d_tasksPV, B.tasksPVCap, maxViews * (size_t)maxWordsPerView) || grow(B.d_pipe, B.pipeCap, kPipeHead + 5 * maxViews))) ||
			    grow(B.d_strips, B.stripCap, maxLines * (size_t)lineRes) ||
			    grow(B.d_lineMask, B.maskCap, maxLines * (size_t)words) || grow(B.d_tabs, B.tabCap, maxViews * tabF))
				return -2;
			if (!B.d_counters)
				OSB_CUDA(cudaMalloc(&B.d_counters, kCounterWords * sizeof(uint32_t)));
		}
		OSB_CUDA(cudaMemcpyAsync(T->d_views, T->h_views, sizeof(ViewParams) * nViews, cudaMemcpyHostToDevice,
		                         c->stream));
		// the pinned view staging is free again once this copy has run: the host prepares the NEXT call's views while
		// this call's kernels execute (waiting for the end of the call here cost 0.45 ms per 4096-view step)
		if (nViews > 8)
			OSB_CUDA(cudaEventRecord(T->stageFree, c->stream));
		if (hostPrep) {
			if (grow(T->h_lineDirs, T->hLineCap, maxLines, true))
				return -2;
			if (grow(T->h_tabs, T->hTabCap, maxViews * tabF, true))
				return -2;
		}
		if (dual) {
			OSB_CUDA(cudaEventRecord(T->fork, c->stream));
			for (ChunkBufs &B : T->set)
				OSB_CUDA(cudaStreamWaitEvent(B.stream, T->fork, 0));
		}
		osbn::SseTables sse{c->d_rcp, c->d_rsqrt, c->rcpBits, c->rsqrtBits};
		const int num = c->opts.numerics;
		const size_t walkSmem = (size_t)(words + kWalkExtraWords) * kWalkThreads * sizeof(uint32_t);
		if (!T->walkBlocksPerSm[num] || T->walkSmem[num] != walkSmem) {
			T->walkSmem[num] = walkSmem;
			if (num == OSB200_NUM_IEEE)
				T->walkBlocksPerSm[num] = resident_blocks(k_walk_lines<false, false>, kWalkThreads, walkSmem);
			else if (num == OSB200_NUM_SSE)
				T->walkBlocksPerSm[num] = resident_blocks(k_walk_lines<true, false>, kWalkThreads, walkSmem);
			else
				T->walkBlocksPerSm[num] = resident_blocks(k_walk_lines<true, true>, kWalkThreads, walkSmem);
		}
		for (size_t ci = 0; ci < nChunks; ci++) {
			ChunkBufs &B = T->set[dual ? (ci & 1) : 0];
			const cudaStream_t st = dual ? B.stream : c->stream;
			const int v0 = chunkStart[ci], nv = chunkStart[ci + 1] - v0;
			const ViewParams &last = T->h_views[v0 + nv - 1];
			const uint32_t totalLines = last.lineBase + (uint32_t)last.numLines;
			const ViewParams *dV = T->d_views + v0;
			int maxViewLines = 0;
			for (int i = 0; i < nv; i++)
				maxViewLines = std::max(maxViewLines, T->h_views[v0 + i].numLines);
			OSB_CUDA(cudaMemsetAsync(B.d_counters, 0, kCounterWords * sizeof(uint32_t), st));
			if (pipeline_window() && B.d_pipe) {
				OSB_CUDA(cudaMemsetAsync(B.d_pipe, 0, (kPipeHead + 4 * (size_t)nv) * sizeof(uint32_t), st));
				OSB_CUDA(cudaMemsetAsync(B.d_pipe + kPipeHead + 4 * (size_t)nv, 0xff, (size_t)nv * sizeof(uint32_t), st));
			}
			if (hostPrep) {
				for (int i = 0; i < nv; i++) {
					const ViewParams &V = T->h_views[v0 + i];
					float *vx = T->h_tabs + V.tabBase, *vy = vx + (w >> 2), *v1 = vy + (h >> 2);
					prep_line_dirs(V, T->h_lineDirs + V.lineBase);
					prep_fog_vx(V, g, vx);
					for (int t = 0; t < g.swThreads; t++) {
						prep_fog_vy(V, g, t, vy);
						prep_v1(V, g, t, v1);
					}
				}
				OSB_CUDA(cudaMemcpyAsync(B.d_lineDirs, T->h_lineDirs, sizeof(float2) * totalLines,
				                         cudaMemcpyHostToDevice, st));
				OSB_CUDA(cudaMemcpyAsync(B.d_tabs, T->h_tabs, sizeof(float) * tabF * nv, cudaMemcpyHostToDevice, st));
			} else {
				KLaunch kl(c, K_VIEW_PREP, st);
				k_view_prep<<<dim3((nv + 31) / 32, 2 + 2 * g.swThreads, 1), 32, 0, st>>>(dV, nv, g, B.d_lineDirs, B.d_tabs);
			}
			// small launches cut a view's lines into smaller tasks (a line is one long dependent chain: what counts is
			// how many warps run side by side and how little each warp diverges)
			int taskLines = 32;
			while (taskLines > line_min_lanes() && (size_t)totalLines / taskLines < (size_t)c->smCount * 4 * 2) // < 2 warps per scheduler
				taskLines >>= 1;
			taskLines = std::max(taskLines, kMinTaskLines);
			// few lines (a single frame, a handful of views): one WARP per line instead of one lane per line
			const bool coop = coop_mode() == 1 || (coop_mode() < 0 && (size_t)totalLines < (size_t)c->smCount * 64);
			if (coop)
				taskLines = 1;
			// batches (tasks of 32 lines, many more tasks than resident warps): tasks queued longest-first
			const bool pipeline = pipeline_window() > 0 && !coop && taskLines == 32 && !hostPrep && !g.noMap;
			const bool ordered = task_order() && !coop && taskLines == 32 && !pipeline;
			const uint32_t bucketCap = ordered ? (uint32_t)(totalLines / 32 + nv) : 0u;
			{
				KLaunch kl(c, K_LINE_SETUP, st);
				const dim3 grid((maxViewLines + kSetupThreads - 1) / kSetupThreads, nv, 1);
				if (num == OSB200_NUM_IEEE)
					k_line_setup<false><<<grid, kSetupThreads, 0, st>>>(dV, B.d_lineDirs, sse, lineRes, g.noMap, B.d_lineTan,
					                                                    B.d_liveBits, maxWordsPerView);
				else
					k_line_setup<true><<<grid, kSetupThreads, 0, st>>>(dV, B.d_lineDirs, sse, lineRes, g.noMap, B.d_lineTan,
					                                                   B.d_liveBits, maxWordsPerView);
				k_line_compact<<<nv, kCompactThreads, 0, st>>>(dV, B.d_lineTan, B.d_liveBits, maxWordsPerView, taskLines,
				                                              B.d_lineRecs, B.d_active, B.d_tasks, B.d_counters, B.d_lineDirs, lineRes,
				                                              ordered ? B.d_bucketTasks : nullptr, bucketCap,
				                                              pipeline ? B.d_tasksPV : nullptr, pipeline ? B.d_pipe + kPipeHead : nullptr,
				                                              maxWordsPerView);
				c->launches++;
			}
			if (pipeline) {
				// walk + gather of a view back to back in one persistent kernel (see k_view_pipeline)
				const int segsP = ((h >> 3) + kSegRows - 1) / kSegRows;
				const uint32_t perView = (uint32_t)((((w >> 3) + 3) >> 2) * segsP);
				WalkIO io{dV, B.d_lineDirs, B.d_lineTan, B.d_lineRecs, c->d_solid, c->d_color, B.d_strips, B.d_lineMask, lineRes};
				GatherIO gioP{dV,      nv,          g, B.d_tabs, c->d_solid, T->d_atan, B.d_strips, B.d_lineRecs, B.d_lineTan, B.d_lineMask,
				              lineRes, d_fb + (size_t)v0 * w * h, d_depth + (size_t)v0 * w * h, kSegRows};
				const unsigned blocks = (unsigned)((size_t)c->smCount * T->walkBlocksPerSm[num]);
				const uint32_t window = (uint32_t)pipeline_window();
				KLaunch kl(c, K_VIEW_PIPELINE, st);
				if (num == OSB200_NUM_IEEE)
					k_view_pipeline<false, false><<<blocks, kWalkThreads, walkSmem, st>>>(io, gioP, sse, B.d_active, B.d_tasksPV, maxWordsPerView,
					                                                                   B.d_pipe, window, perView, (uint32_t)pipeline_split());
				else if (num == OSB200_NUM_SSE)
					k_view_pipeline<true, false><<<blocks, kWalkThreads, walkSmem, st>>>(io, gioP, sse, B.d_active, B.d_tasksPV, maxWordsPerView,
					                                                                  B.d_pipe, window, perView, (uint32_t)pipeline_split());
				else
					k_view_pipeline<true, true><<<blocks, kWalkThreads, walkSmem, st>>>(io, gioP, sse, B.d_active, B.d_tasksPV, maxWordsPerView,
					                                                                 B.d_pipe, window, perView, (uint32_t)pipeline_split());
#ifdef OSB_PIPE_STATS
				{
					unsigned long long hs[4];
					cudaMemcpyAsync(hs, B.d_pipe + 96, sizeof(hs), cudaMemcpyDeviceToHost, st);
					cudaStreamSynchronize(st);
					fprintf(stderr, "pipe stats: blocks %u  dispatch %.3f  walk %.3f  gather %.3f Gcycles (lane 0 of every warp), sleeps %llu\n", blocks,
					        hs[0] * 1e-9, hs[1] * 1e-9, hs[2] * 1e-9, hs[3]);
				}
#endif
			} else if (!g.noMap && coop) {
				WalkIO io{dV, B.d_lineDirs, B.d_lineTan, B.d_lineRecs, c->d_solid, c->d_color, B.d_strips, B.d_lineMask, lineRes};
				const size_t smem = coop_shared_bytes(lineRes);
				const unsigned blocks = (unsigned)std::min<size_t>(totalLines, (size_t)c->smCount * 32);
				KLaunch kl(c, K_WALK_LINES, st);
				if (num == OSB200_NUM_IEEE)
					k_walk_coop<false, false><<<blocks, 32, smem, st>>>(io, sse, B.d_active, B.d_tasks, B.d_counters);
				else if (num == OSB200_NUM_SSE)
					k_walk_coop<true, false><<<blocks, 32, smem, st>>>(io, sse, B.d_active, B.d_tasks, B.d_counters);
				else
					k_walk_coop<true, true><<<blocks, 32, smem, st>>>(io, sse, B.d_active, B.d_tasks, B.d_counters);
			} else if (!g.noMap) {
				const size_t wantWarps = ((size_t)totalLines + taskLines - 1) / taskLines;
				int perSm = T->walkBlocksPerSm[num];
				if (walk_blocks_cap() && nChunks > 1)
					perSm = std::min(perSm, walk_blocks_cap());
				const size_t resident = (size_t)c->smCount * perSm;
				const unsigned blocks = (unsigned)std::min<size_t>((wantWarps + kWalkThreads / 32 - 1) / (kWalkThreads / 32), resident);
				WalkIO io{dV, B.d_lineDirs, B.d_lineTan, B.d_lineRecs, c->d_solid, c->d_color, B.d_strips, B.d_lineMask, lineRes};
				KLaunch kl(c, K_WALK_LINES, st);
#define LAUNCH_WALK(SSE, SH)                                                                                                      \
	do {                                                                                                                      \
		if (SSE && sse.rcpBits <= kRcpSharedBits && rcp_shared())                                                              \
			k_walk_lines<SSE, SH, SSE><<<blocks, kWalkThreads, walkSmem, st>>>(io, sse, B.d_active,                              \
			                                                                 ordered ? B.d_bucketTasks : B.d_tasks, B.d_counters, \
			                                                                 bucketCap);                                          \
		else                                                                                                                  \
			k_walk_lines<SSE, SH><<<blocks, kWalkThreads, walkSmem, st>>>(io, sse, B.d_active,                                   \
			                                                              ordered ? B.d_bucketTasks : B.d_tasks, B.d_counters,    \
			                                                              bucketCap);                                             \
	} while (0)
				if (num == OSB200_NUM_IEEE)
					LAUNCH_WALK(false, false);
				else if (num == OSB200_NUM_SSE)
					LAUNCH_WALK(true, false);
				else
					LAUNCH_WALK(true, true);
#undef LAUNCH_WALK
			}
			// warp tasks of the gather: 8 block rows each for batches; a single frame is cut finer (2 block rows) so that
			// its few hundred tasks become enough warps to hide the gather's load latency
			const size_t colGroupsAll = (size_t)nv * (((w >> 3) + 3) >> 2);
			int segRows = kSegRows;
			while (segRows > 2 && colGroupsAll * (((h >> 3) + segRows - 1) / segRows) < (size_t)c->smCount * 32)
				segRows >>= 1;
			const int segs = ((h >> 3) + segRows - 1) / segRows;
			const size_t tasks = colGroupsAll * segs;
			const unsigned gBlocks = (unsigned)((tasks + kGatherWarps - 1) / kGatherWarps);
			GatherIO gio{dV,           nv,      g, B.d_tabs, c->d_solid, T->d_atan, B.d_strips, B.d_lineRecs, B.d_lineTan, B.d_lineMask,
			             lineRes, d_fb + (size_t)v0 * w * h, d_depth + (size_t)v0 * w * h, segRows};
			if (!pipeline) {
				KLaunch kl(c, K_GATHER_FOG, st);
				if (num == OSB200_NUM_IEEE)
					k_gather_fog<false, false><<<gBlocks, kGatherWarps * 32, 0, st>>>(gio, sse);
				else if (num == OSB200_NUM_SSE)
					k_gather_fog<true, false><<<gBlocks, kGatherWarps * 32, 0, st>>>(gio, sse);
				else
					k_gather_fog<true, true><<<gBlocks, kGatherWarps * 32, 0, st>>>(gio, sse);
			}
			OSB_CUDA(cudaEventRecord(B.done, st));
			if (host_fb || host_depth) {
				const size_t off = (size_t)v0 * w * h, cnt = (size_t)nv * w * h;
				OSB_CUDA(cudaStreamWaitEvent(T->copyStream, B.done, 0));
				if (host_fb)
					OSB_CUDA(cudaMemcpyAsync(host_fb + off, d_fb + off, cnt * 4, cudaMemcpyDeviceToHost, T->copyStream));
				if (host_depth)
					OSB_CUDA(cudaMemcpyAsync(host_depth + off, d_depth + off, cnt * 4, cudaMemcpyDeviceToHost, T->copyStream));
			}
			if (hostPrep && ci + 1 < nChunks) // staging reused by the next chunk
				OSB_CUDA(cudaStreamSynchronize(st));
		}
		if (dual) { // join: the context stream continues after both working streams
			for (ChunkBufs &B : T->set)
				OSB_CUDA(cudaStreamWaitEvent(c->stream, B.done, 0));
		}
		OSB_CUDA(cudaGetLastError());
		if (hostPrep) // line directions / tables were staged in pinned memory as well
			OSB_CUDA(cudaEventRecord(T->stageFree, c->stream));
		return 0;
	}
	// Context-owned frame + depth of the host-pointer entry points.  The depth buffer persists between frames like
	// SWRenderer::depthBuffer (never cleared, SWRenderer.cpp:83); a fresh one starts at zero like vector::resize.
	int terrain_out_buffers(osb200_ctx *c, size_t n, uint32_t **fb, float **depth) {
		if (terrain_init(c))
			return -2;
		TerrainState *T = c->terrain;
		if (n != T->outCap) {
			cudaFree(T->d_fb);
			cudaFree(T->d_depth);
			T->d_fb = nullptr;
			T->d_depth = nullptr;
			T->outCap = 0;
			OSB_CUDA(cudaMalloc(&T->d_fb, n * 4));
			OSB_CUDA(cudaMalloc(&T->d_depth, n * 4));
			OSB_CUDA(cudaMemsetAsync(T->d_depth, 0, n * 4, c->stream));
			T->outCap = n;
		}
		// the previous call's copies must not race with this call's kernels overwriting the same device buffers
		OSB_CUDA(cudaStreamSynchronize(T->copyStream));
		*fb = T->d_fb;
		*depth = T->d_depth;
		return 0;
	}

	// ApplyFog over the single view last set up by render_impl(.., deferFog = true) on the context stream
	// The tail of SWRenderer::EndScene for a frame that drew models and / or dynamic lights, in one pass over the frame:
	//   winner of the splat keys (SWModelRenderer's depth-tested writes, models.cu)      keys == null: nothing was splatted
	//   ApplyDynamicLight (SWRenderer.cpp:283-325) for every kept light in submission order, float recurrences from
	//   the tables of k_light_prep                                                       nLights == 0: no lights
	//   ApplyFog (SWRenderer.cpp:341-516), same per-4x4-block scale and blend as k_gather_fog
	template <bool kSse, bool kSse2Fog>
	__global__ void __launch_bounds__(256)
	  k_scene_finish(const ViewParams *__restrict__ views, FrameGeom g, const float *__restrict__ tabs, osbn::SseTables sse,
	                 uint32_t *__restrict__ fb, float *__restrict__ depthBuf, const unsigned long long *__restrict__ keys,
	                 const uint32_t *__restrict__ splatColor, const LightDev *__restrict__ lights, int nLights,
	                 const float *__restrict__ lightTabs) {
		const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
		if (x >= g.fw || y >= g.fh)
			return;
		const size_t o = (size_t)y * g.fw + x;
		uint32_t color;
		float z;
		bool splat = false;
		if (keys) {
			const unsigned long long k = keys[o];
			const uint32_t seq = (uint32_t)k;
			if (seq) {
				color = splatColor[seq];
				z = key_depth((uint32_t)(k >> 32));
				depthBuf[o] = z;
				splat = true;
			}
		}
		if (!splat) {
			color = fb[o];
			z = depthBuf[o];
		}
		for (int li = 0; li < nLights; li++) {
			const LightDev &L = lights[li];
			if (x < L.minX || x >= L.maxX || y < L.minY || y >= L.maxY)
				continue;
			const float *vxT = lightTabs + L.tabBase, *vyT = vxT + g.fw;
			const float px = __fsub_rn(__fmul_rn(vxT[x], z), L.center[0]);
			const float py = __fsub_rn(__fmul_rn(vyT[y], z), L.center[1]);
			const float pz = __fsub_rn(z, L.center[2]);
			float dist = __fadd_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)), __fmul_rn(pz, pz));
			dist = __fmul_rn(dist, L.invRadius2);
			if (dist < 1.f) {
				float strength = __fsub_rn(1.f, dist);
				strength = __fmul_rn(strength, strength);
				strength = __fmul_rn(strength, 256.f);
				const int factor = osbn::x86_f2i(strength);
				const uint32_t sr = (color >> 16) & 0xff, sg = (color >> 8) & 0xff, sb = color & 0xff;
				const int ar = (int)((uint32_t)(L.r * factor) * sr), ag = (int)((uint32_t)(L.g * factor) * sg),
				          ab = (int)((uint32_t)(L.b * factor) * sb);
				const uint32_t dr = min((uint32_t)(ar >> 16) + sr, 255u), dg = min((uint32_t)(ag >> 16) + sg, 255u),
				               db = min((uint32_t)(ab >> 16) + sb, 255u);
				color = db | (dg << 8) | (dr << 16);
			}
		}
		const ViewParams &V = views[0];
		osbn::Num<kSse> num{sse};
		const float *vxTab = tabs + V.tabBase, *vyTab = vxTab + (g.fw >> 2);
		const float vx = vxTab[x >> 2], vy = vyTab[y >> 2];
		float ds = __fadd_rn(__fadd_rn(1.f, __fmul_rn(vx, vx)), __fmul_rn(vy, vy));
		ds = __fmul_rn(ds, __fmul_rn(num.rsqrt(ds), V.fogScale));
		const uint32_t fogBR = g.fogB | (g.fogR << 16), fogG8 = g.fogG << 8;
		fb[o] = fog_blend<kSse2Fog>(color, __fmul_rn(z, ds), fogBR, fogG8);
	}

	int terrain_finish_scene(osb200_ctx *c, int w, int h, const float fogColor[3], uint32_t *d_fb, float *d_depth,
	                         const unsigned long long *d_keys, const uint32_t *d_splatColor, const LightDev *d_lights,
	                         int nLights, const float *d_lightTabs) {
		TerrainState *T = c->terrain;
		FrameGeom g;
		std::memset(&g, 0, sizeof(g));
		g.fw = w;
		g.fh = h;
		g.under = c->opts.undersampling;
		g.swThreads = c->opts.sw_threads;
		g.sse2Fog = g.sse2Shade = c->opts.numerics == OSB200_NUM_SSE2;
		auto toFixed8 = [](float v) {
			int i = static_cast<int>(v * 255.f + .5f);
			return (uint32_t)std::max(std::min(i, 255), 0);
		};
		g.fogR = toFixed8(fogColor[0]);
		g.fogG = toFixed8(fogColor[1]);
		g.fogB = toFixed8(fogColor[2]);
		osbn::SseTables sse{c->d_rcp, c->d_rsqrt, c->rcpBits, c->rsqrtBits};
		const dim3 grid((w + 31) / 32, (h + 7) / 8, 1);
		KLaunch kl(c, K_SCENE_FINISH);
#define LAUNCH_FINISH(SSE, FOG)                                                                                        \
	k_scene_finish<SSE, FOG><<<grid, 256, 0, c->stream>>>(T->d_views, g, T->set[0].d_tabs, sse, d_fb, d_depth, d_keys,   \
	                                                      d_splatColor, d_lights, nLights, d_lightTabs)
		if (c->opts.numerics == OSB200_NUM_IEEE)
			LAUNCH_FINISH(false, false);
		else if (c->opts.numerics == OSB200_NUM_SSE)
			LAUNCH_FINISH(true, false);
		else
			LAUNCH_FINISH(true, true);
#undef LAUNCH_FINISH
		OSB_CUDA(cudaGetLastError());
		return 0;
	}

} // namespace osb

extern "C" {
#ifdef OSB_STATS
__attribute__((visibility("default"))) int osb200_debug_stats(unsigned long long *out, int reset) {
	cudaDeviceSynchronize();
	cudaMemcpyFromSymbol(out, osb::g_stats, sizeof(unsigned long long) * 16);
	if (reset) {
		unsigned long long z[16] = {0};
		cudaMemcpyToSymbol(osb::g_stats, z, sizeof(z));
	}
	return 0;
}
#endif

int osb200_render_device(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h,
                         const float fogColor[3], float fogDist, uint32_t *d_bgrx, float *d_depth) {
	return osb::render_impl(c, scenes, nViews, w, h, fogColor, fogDist, d_bgrx, d_depth, nullptr, nullptr, false);
}

int osb200_render(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
                  float fogDist, uint32_t *host_bgrx, float *host_depth) {
	if (!c) {
		osb::set_error("osb200_render: no context");
		return -1;
	}
	if (nViews <= 0)
		return nViews == 0 ? 0 : -1;
	if (w <= 0 || h <= 0 || (w & 7) || (h & 7)) {
		osb::set_error("Framebuffer size is not multiple of 8.");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	if (osb::terrain_init(c))
		return -2;
	osb::TerrainState *T = c->terrain;
	uint32_t *ofb;
	float *odp;
	if (int rc0 = osb::terrain_out_buffers(c, (size_t)nViews * w * h, &ofb, &odp))
		return rc0;
	int rc = osb::render_impl(c, scenes, nViews, w, h, fogColor, fogDist, T->d_fb, T->d_depth, host_bgrx, host_depth, false);
	if (rc)
		return rc;
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	OSB_CUDA(cudaStreamSynchronize(T->copyStream));
	return 0;
}

} // extern "C"
namespace osb {
	// group.cu: render into the context-owned device buffers and hand them out (no host copy, no sync)
	int render_device_frames(osb200_ctx *c, const osb200_scene *scenes, int nViews, int w, int h, const float fogColor[3],
	                         float fogDist, uint32_t **d_fb, float **d_depth) {
		if (!c) {
			set_error("osb200_render: no context");
			return -1;
		}
		OSB_CUDA(cudaSetDevice(c->device));
		if (int rc = terrain_out_buffers(c, (size_t)nViews * w * h, d_fb, d_depth))
			return rc;
		return render_impl(c, scenes, nViews, w, h, fogColor, fogDist, *d_fb, *d_depth, nullptr, nullptr, false);
	}
} // namespace osb
extern "C" {

int osb200_read_depth(osb200_ctx *c, float *out, size_t n) {
	if (!c || !out) {
		osb::set_error("osb200_read_depth: null argument");
		return -1;
	}
	if (!c->terrain || n > c->terrain->outCap) {
		osb::set_error("osb200_read_depth: no frame of that size has been rendered by osb200_render");
		return -1;
	}
	OSB_CUDA(cudaSetDevice(c->device));
	OSB_CUDA(cudaMemcpyAsync(out, c->terrain->d_depth, n * 4, cudaMemcpyDeviceToHost, c->stream));
	OSB_CUDA(cudaStreamSynchronize(c->stream));
	return 0;
}

} // extern "C"
