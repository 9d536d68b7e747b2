// Multi-device handle: base-set sharding behind the C ABI (SURVEY.md §8(b) "cfg carries the device list, the handle
// owns all device memory and streams", §8(e)). One process, one expann_gpu_t, G GPUs.
//
// The reference engines live in one address space and are driven by one thread (src/ann_engine.h:16-29,
// src/bench_runner.h:61-87, src/basic_bench.h:65-84), so a drop-in engine that wants more than one GPU has to fan out
// below the engine interface. A multi-device handle owns one ordinary single-device handle ("kid") per GPU:
//
//   decomposition  G = gq x gr: the rows are cut into gr contiguous shards (rows [ir*ceil(n/gr), ...), global id =
//                  local id + offset, added on the device), each shard is held by gq GPUs, and query group iq of a call
//                  (a contiguous block of the m queries) is served by the gr kids (iq, *). gr = G is the plain row
//                  sharding of §8(e) ("mgpu_row_groups" = G); by default a shard holds up to 2^20 rows and the GPUs that
//                  are not needed to hold the base split the queries (the tensor-core filter warms a bound up per row
//                  stream: slivers of a small base cost more than they save); a graph index read from a file is
//                  replicated (gr = 1: queries split, no merge).
//   query fan-out  the queries reach GPU 0 once (one H2D copy, or the caller's device buffer); every other kid pulls ITS
//                  block over NVLink (cudaMemcpyPeerAsync on its own stream) — not G copies over PCIe.
//   fused tail     the last kernel of a kid's path (K4 rerank / the fused split merge of K1 / the graph search) stores
//                  its [m][k] keys straight into GPU 0's gather buffer through the peer mapping (NVLink stores issued by
//                  the compute kernel itself: there is no gather step, no NCCL call, no copy kernel); GPU 0's stream waits
//                  on the kids' events and K7 merges the gr lists of every query. With gr = 1 the kids write the final
//                  [m][k] block and nothing is merged. The gather buffer is double-buffered by call parity, so the kids
//                  of call s+1 overlap GPU 0's merge of call s.
//   result         keys carry (key, id), so the merged result is bit-identical to the one-GPU result whatever G, gq, gr.
// Every kid's launches are issued by its own parked host thread (kid 0: the calling thread); every launch is asynchronous,
// cross-device ordering is by CUDA events.
#include "engine.h"
#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <utility>

namespace expann {

// One parked host thread per kid: a call's per-kid launches (a dozen driver calls each) are issued side by side instead
// of one kid after the other, so the last GPU starts its filter microseconds, not G x 30 us, after the first.
struct MgWorker {
	std::thread th;
	std::mutex mu;
	std::condition_variable cv;
	std::function<int()> task;
	bool has_task = false, done = true, quit = false;
	int rc = 0;
	void start(int device) {
		th = std::thread([this, device] {
			cudaSetDevice(device);
			cudaFree(nullptr); // bind the primary context: driver entry points (cuTensorMapEncodeTiled) need one
			std::unique_lock<std::mutex> lk(mu);
			for (;;) {
				cv.wait(lk, [this] { return has_task || quit; });
				if (quit) return;
				has_task = false;
				lk.unlock();
				const int r = task();
				lk.lock();
				rc = r;
				done = true;
				cv.notify_all();
			}
		});
	}
	void post(std::function<int()> f) {
		std::lock_guard<std::mutex> lk(mu);
		task = std::move(f);
		has_task = true;
		done = false;
		cv.notify_all();
	}
	int wait() {
		std::unique_lock<std::mutex> lk(mu);
		cv.wait(lk, [this] { return done; });
		return rc;
	}
	void stop() {
		{
			std::lock_guard<std::mutex> lk(mu);
			quit = true;
			cv.notify_all();
		}
		if (th.joinable()) th.join();
	}
};

struct Multi {
	std::vector<int> devices;        // one entry per kid slot; an ordinal may repeat (several shards on one GPU: tests)
	std::vector<expann_gpu*> kids;   // kid (iq, ir) at [iq * gr + ir]
	int gq = 1, gr = 1;
	int opt_row_groups = 0;          // 0 = automatic (shards of up to MG_SHARD_ROWS rows), R = R row shards
	std::vector<float> rows;         // host copy of the stored rows until the shards are uploaded (store_vector copies too)
	std::vector<std::pair<std::string, int64_t>> options; // replayed on the kids when they are created
	std::string index_path;          // graph engine: every kid reads the same index file (replicas)
	bool materialized = false, timing = false;
	size_t per_rows = 0;
	// GPU 0 side
	DevBuf gather[2], q_dev, keys, res;
	cudaEvent_t ev_q = nullptr, ev_merged[2] = { nullptr, nullptr };
	bool merged_pending[2] = { false, false };
	// per kid, twice (index 2*kid + parity of the call): consecutive calls alternate, so that a brute-force kid can work on
	// two calls side by side (its second stream + the handle's second set of work buffers, api.cu scratch_for_stream)
	std::vector<cudaEvent_t> ev_done;
	std::vector<DevBuf> kid_q, kid_keys;
	std::vector<cudaStream_t> kid_stream2; // second stream of a brute-force kid (nullptr: one stream, calls run in order)
	std::vector<char> direct, same_dev;
	uint64_t step = 0, peer_bytes = 0;
	std::vector<std::unique_ptr<MgWorker>> workers; // [kid]; kid 0's work runs on the calling thread
	std::vector<std::string> kid_err;
};

constexpr size_t MG_SHARD_ROWS = (size_t)1 << 20; // automatic decomposition: rows per shard before another shard is opened

static size_t round_up_sz(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int kid_fail(expann_gpu* h, expann_gpu* kid, int rc, size_t c) {
	return fail(h, rc, "GPU shard " + std::to_string(c) + " (device " + std::to_string(kid->device) + "): " + kid->err);
}

int mg_create(const expann_gpu_cfg* cfg, const int32_t* devices, size_t ndev, expann_gpu** out) {
	*out = nullptr;
	if (!devices || ndev == 0 || ndev > 64) return fail(nullptr, EXPANN_ERR_ARG, "create_multi: 1..64 devices");
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (e != cudaSuccess || count == 0)
		return fail(nullptr, EXPANN_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this library has no CPU fallback)");
	for (size_t i = 0; i < ndev; ++i) {
		if (devices[i] < 0 || devices[i] >= count) return fail(nullptr, EXPANN_ERR_ARG, "create_multi: device ordinal out of range");
		cudaDeviceProp prop;
		e = cudaGetDeviceProperties(&prop, devices[i]);
		if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceProperties");
		if (prop.major != 10) return fail(nullptr, EXPANN_ERR_UNSUPPORTED, "device is not sm_100 (kernels are built for sm_100a only)");
	}
	expann_gpu* h = new (std::nothrow) expann_gpu();
	if (!h) return fail(nullptr, EXPANN_ERR_ARG, "out of host memory");
	h->cfg = *cfg;
	h->quantizer = cfg->quantizer;
	h->device = devices[0];
	h->mg = new Multi();
	h->mg->devices.assign(devices, devices + ndev);
	h->mg->opt_row_groups = cfg->row_groups > 0 ? cfg->row_groups : 0;
	DeviceGuard g(h->device);
	e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
	if (e != cudaSuccess) { delete h->mg; delete h; return cuda_fail(nullptr, e, "cudaStreamCreate"); }
	*out = h;
	return EXPANN_OK;
}

void mg_destroy(expann_gpu* h) {
	Multi* g = h->mg;
	if (!g) return;
	for (auto& w : g->workers)
		if (w) w->stop();
	g->workers.clear();
	for (expann_gpu* kid : g->kids)
		if (kid) { DeviceGuard kg(kid->device); cudaStreamSynchronize(kid->stream); }
	for (size_t c = 0; c < g->kid_stream2.size(); ++c)
		if (g->kid_stream2[c] && g->kids[c]) { DeviceGuard kg(g->kids[c]->device); cudaStreamSynchronize(g->kid_stream2[c]); }
	for (size_t c = 0; c < g->kids.size(); ++c) {
		if (!g->kids[c]) continue;
		{
			DeviceGuard kg(g->kids[c]->device);
			for (size_t j = 2 * c; j < 2 * c + 2; ++j) {
				if (j < g->ev_done.size() && g->ev_done[j]) cudaEventDestroy(g->ev_done[j]);
				if (j < g->kid_q.size()) g->kid_q[j].release();
				if (j < g->kid_keys.size()) g->kid_keys[j].release();
			}
			if (c < g->kid_stream2.size() && g->kid_stream2[c]) cudaStreamDestroy(g->kid_stream2[c]);
		}
		expann_gpu_destroy(g->kids[c]);
	}
	DeviceGuard rg(h->device);
	if (h->stream) cudaStreamSynchronize(h->stream);
	if (g->ev_q) cudaEventDestroy(g->ev_q);
	for (int p = 0; p < 2; ++p) {
		if (g->ev_merged[p]) cudaEventDestroy(g->ev_merged[p]);
		g->gather[p].release();
	}
	g->q_dev.release();
	g->keys.release();
	g->res.release();
	delete g;
	h->mg = nullptr;
}

int mg_add(expann_gpu* h, const float* rows, size_t n, size_t d, bool device_rows, cudaStream_t st) {
	Multi& g = *h->mg;
	if (d == 0) return fail(h, EXPANN_ERR_ARG, "d must be > 0");
	if (g.materialized) return fail(h, EXPANN_ERR_STATE, "multi-device handle: rows cannot be added once the shards are on the GPUs");
	if (!g.index_path.empty()) return EXPANN_OK; // read_index: stored vectors are ignored (src/antitopo_engine.h:312-313)
	if (h->d == 0) { h->d = d; h->d_pad = round_up_sz(d, 32); }
	else if (h->d != d) return fail(h, EXPANN_ERR_ARG, "dimension mismatch with previously stored vectors");
	if (n == 0) return EXPANN_OK;
	try {
		const size_t old = g.rows.size();
		g.rows.resize(old + n * d);
		if (device_rows) {
			DeviceGuard rg(h->device);
			EXPANN_CUDA(h, cudaStreamSynchronize(st));
			EXPANN_CUDA(h, cudaMemcpy(g.rows.data() + old, rows, n * d * sizeof(float), cudaMemcpyDeviceToHost));
		} else
			std::memcpy(g.rows.data() + old, rows, n * d * sizeof(float));
	} catch (const std::exception&) {
		return fail(h, EXPANN_ERR_ARG, "out of host memory");
	}
	h->n += n;
	h->built = false;
	return EXPANN_OK;
}

// choose gq x gr, create the kids, upload the shards (or have every kid read the index file), open the peer mappings
static int mg_materialize(expann_gpu* h) {
	Multi& g = *h->mg;
	if (g.materialized) return EXPANN_OK;
	const size_t G = g.devices.size();
	const bool replicas = !g.index_path.empty();
	if (!replicas && h->n == 0) return fail(h, EXPANN_ERR_STATE, "build() on an empty engine (the reference asserts n > 0)");
	// Row shards. Explicit ("mgpu_row_groups" = R): R shards. Automatic (0): shards of up to MG_SHARD_ROWS rows -- a base that
	// is small against the GPU count is NOT cut into slivers: the tensor-core filter warms a bound up per row stream, and on
	// a 125k-row shard (C2 over 8 GPUs) that warm-up is half the kernel time, so the GPUs that are not needed to hold the rows
	// split the QUERIES of a call instead (same rows on each GPU of a group; measured, profiles/r2_notes.md). C4 (10M rows) is
	// cut into one shard per GPU either way.
	size_t gr;
	if (replicas) gr = 1;
	else if (g.opt_row_groups > 0) gr = std::min<size_t>(G, (size_t)g.opt_row_groups);
	else {
		gr = std::min<size_t>(G, std::max<size_t>(1, (h->n + MG_SHARD_ROWS - 1) / MG_SHARD_ROWS));
		while (G % gr) ++gr; // the next count of shards that leaves whole query groups
	}
	while (G % gr) --gr; // whole query groups
	size_t per = 0;
	if (!replicas) {
		per = (h->n + gr - 1) / gr;
		gr = (h->n + per - 1) / per; // no empty shards (tiny bases)
	}
	const size_t gq = std::max<size_t>(1, G / gr);
	g.gr = (int)gr;
	g.gq = (int)gq;
	g.per_rows = per;
	const size_t nk = gq * gr;
	g.kids.assign(nk, nullptr);
	g.ev_done.assign(2 * nk, nullptr);
	g.kid_q.resize(2 * nk);
	g.kid_keys.resize(2 * nk);
	g.kid_stream2.assign(nk, nullptr);
	g.direct.assign(nk, 1);
	g.same_dev.assign(nk, 0);
	for (size_t c = 0; c < nk; ++c) {
		const size_t ir = c % gr;
		expann_gpu_cfg kc = h->cfg;
		kc.device = g.devices[c];
		kc.n_devices = 0;
		kc.row_groups = 0;
		const size_t lo = replicas ? 0 : ir * per, hi = replicas ? 0 : std::min(h->n, lo + per);
		kc.id_offset = h->cfg.id_offset + lo;
		expann_gpu* kid = nullptr;
		if (int rc = expann_gpu_create(&kc, &kid)) return fail(h, rc, std::string("GPU shard create: ") + expann_gpu_last_error(nullptr));
		g.kids[c] = kid;
		for (auto& o : g.options)
			if (int rc = expann_gpu_set_option(kid, o.first.c_str(), o.second)) return kid_fail(h, kid, rc, c);
		if (g.timing) expann_gpu_set_timing(kid, 1);
		if (replicas) {
			if (int rc = expann_gpu_load_index(kid, g.index_path.c_str())) return kid_fail(h, kid, rc, c);
		} else if (int rc = expann_gpu_add(kid, g.rows.data() + lo * h->d, hi - lo, h->d))
			return kid_fail(h, kid, rc, c);
		g.same_dev[c] = kid->device == h->device;
		DeviceGuard kg(kid->device);
		EXPANN_CUDA(h, cudaEventCreateWithFlags(&g.ev_done[2 * c], cudaEventDisableTiming));
		EXPANN_CUDA(h, cudaEventCreateWithFlags(&g.ev_done[2 * c + 1], cudaEventDisableTiming));
		if (h->cfg.engine == EXPANN_ENGINE_BRUTE_FORCE) EXPANN_CUDA(h, cudaStreamCreateWithFlags(&g.kid_stream2[c], cudaStreamNonBlocking));
		if (!g.same_dev[c]) {
			// the kid's kernels store their keys into GPU 0's gather buffer: needs the peer mapping kid -> GPU 0
			int can = 0;
			EXPANN_CUDA(h, cudaDeviceCanAccessPeer(&can, kid->device, h->device));
			if (can) {
				cudaError_t e = cudaDeviceEnablePeerAccess(h->device, 0);
				if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
				else if (e != cudaSuccess) can = 0, cudaGetLastError();
			}
			g.direct[c] = (char)can; // without it the keys go back with a peer copy behind the kid's last kernel
			int back = 0;
			DeviceGuard rg(h->device);
			if (cudaDeviceCanAccessPeer(&back, h->device, kid->device) == cudaSuccess && back) {
				cudaError_t e = cudaDeviceEnablePeerAccess(kid->device, 0);
				if (e != cudaSuccess) cudaGetLastError();
			}
		}
	}
	if (replicas) {
		h->n = g.kids[0]->n;
		h->d = g.kids[0]->d;
		h->d_pad = g.kids[0]->d_pad;
	}
	DeviceGuard rg(h->device);
	EXPANN_CUDA(h, cudaEventCreateWithFlags(&g.ev_q, cudaEventDisableTiming));
	for (int p = 0; p < 2; ++p) EXPANN_CUDA(h, cudaEventCreateWithFlags(&g.ev_merged[p], cudaEventDisableTiming));
	g.workers.resize(nk);
	g.kid_err.assign(nk, std::string());
	for (size_t c = 1; c < nk; ++c) {
		g.workers[c].reset(new MgWorker());
		g.workers[c]->start(g.kids[c]->device);
	}
	g.materialized = true;
	return EXPANN_OK;
}

// run f on every kid, one host thread per kid (the builds are seconds of per-GPU work with host synchronisations inside:
// side by side they cost one GPU's time, not G times that); first failure wins
template <class F> static int mg_each_kid(expann_gpu* h, F f) {
	Multi& g = *h->mg;
	std::vector<int> rcs(g.kids.size(), 0);
	std::vector<std::thread> th;
	for (size_t c = 1; c < g.kids.size(); ++c)
		th.emplace_back([&rcs, &g, &f, c] {
			// a new host thread has no CUDA context bound: driver entry points (cuTensorMapEncodeTiled) fail with
			// CUDA_ERROR_INVALID_CONTEXT until a runtime call has made the device's primary context current
			cudaSetDevice(g.kids[c]->device);
			cudaFree(nullptr);
			rcs[c] = f(g.kids[c]);
		});
	if (!g.kids.empty()) rcs[0] = f(g.kids[0]);
	for (auto& t : th) t.join();
	for (size_t c = 0; c < g.kids.size(); ++c)
		if (rcs[c]) return kid_fail(h, g.kids[c], rcs[c], c);
	return EXPANN_OK;
}

int mg_load_index(expann_gpu* h, const char* path) {
	Multi& g = *h->mg;
	if (g.materialized) return fail(h, EXPANN_ERR_STATE, "multi-device handle: load_index after the shards were created");
	g.index_path = path;
	g.rows.clear();
	g.rows.shrink_to_fit();
	h->n = 0;
	h->d = 0;
	h->built = false;
	return mg_materialize(h); // every GPU reads the file now (errors surface here, like the single-device call)
}

int mg_build_graph(expann_gpu* h, uint64_t M, uint64_t M0, uint64_t K, float ortho_factor, float ortho_bias, uint64_t prune_overflow,
                   uint64_t seed) {
	Multi& g = *h->mg;
	if (!g.index_path.empty()) return fail(h, EXPANN_ERR_STATE, "build_graph: an index file was loaded");
	if (int rc = mg_materialize(h)) return rc;
	// one sub-graph per row shard (mode (B) of SURVEY 8(e)): every query searches every shard with the same ef
	return mg_each_kid(h, [=](expann_gpu* kid) { return expann_gpu_build_graph(kid, M, M0, K, ortho_factor, ortho_bias, prune_overflow, seed); });
}

int mg_build(expann_gpu* h) {
	Multi& g = *h->mg;
	if (int rc = mg_materialize(h)) return rc;
	if (h->n + h->cfg.id_offset > 0xFFFFFFFEull) return fail(h, EXPANN_ERR_UNSUPPORTED, "ids must fit 32 bits on device");
	if (h->cfg.engine == EXPANN_ENGINE_QUANT_SCAN && h->quantizer == 1 && !g.rows.empty()) {
		// quantizer_ranged_q8::build fits ONE range over all components of all vectors (src/quantizer.h:206-227): fit it over
		// the whole base here, so that the codes (and the result) do not depend on how the rows are sharded
		float mx = 1.17549435e-38f, mn = 3.402823466e+38f;
		for (float v : g.rows) { mx = v > mx ? v : mx; mn = v < mn ? v : mn; }
		for (expann_gpu* kid : g.kids) quant_set_range(kid, mn, mx);
		quant_set_range(h, mn, mx);
	}
	if (int rc = mg_each_kid(h, [](expann_gpu* kid) { return expann_gpu_build(kid); })) return rc;
	g.rows.clear();
	g.rows.shrink_to_fit();
	h->stats.n = h->n;
	h->stats.d = h->d;
	h->stats.d_pad = h->d_pad;
	h->built = true;
	return EXPANN_OK;
}

// queries [q_lo, q_hi) of query group iq
static void mg_query_block(const Multi& g, size_t m, size_t iq, size_t* q_lo, size_t* q_hi) {
	size_t per = (m + g.gq - 1) / g.gq;
	if (m >= (size_t)512 * g.gq) per = round_up_sz(per, 256); // whole 256-query tiles of the tensor-core filter
	*q_lo = std::min(m, iq * per);
	*q_hi = std::min(m, *q_lo + per);
}

// d_q: the queries on GPU 0 (kids on other GPUs pull their block over NVLink), or -- host_q -- in page-locked host memory:
// every kid then copies ITS block straight from the host on its own stream (with query groups each GPU needs only 1/gq of
// the queries, and every GPU has its own PCIe link: G small copies side by side instead of one large one plus a fan-out)
static int mg_query_impl(expann_gpu* h, const float* d_q, const float* host_q, size_t m, size_t k, size_t ef, uint64_t* d_keys_out, cudaStream_t st) {
	Multi& g = *h->mg;
	if (!h->built) return fail(h, EXPANN_ERR_STATE, "query before build()");
	if (k == 0) return fail(h, EXPANN_ERR_ARG, "k must be > 0");
	if (m == 0) return EXPANN_OK;
	DeviceGuard rg(h->device);
	const int par = (int)(g.step & 1);
	++g.step;
	const size_t d = h->d, gr = (size_t)g.gr;
	uint64_t* gat = d_keys_out; // one row group: the kids write the final block themselves
	if (gr > 1) {
		if (g.gather[par].cap < gr * m * k * 8) {
			EXPANN_CUDA(h, cudaDeviceSynchronize()); // growing: nobody may still be writing into the old buffer
			for (size_t c = 0; c < g.kids.size(); ++c) {
				DeviceGuard kg(g.kids[c]->device);
				EXPANN_CUDA(h, cudaStreamSynchronize(g.kids[c]->stream));
				if (g.kid_stream2[c]) EXPANN_CUDA(h, cudaStreamSynchronize(g.kid_stream2[c]));
			}
			EXPANN_CUDA(h, g.gather[par].ensure(gr * m * k * 8));
			g.merged_pending[par] = false;
		}
		gat = g.gather[par].as<uint64_t>();
	}
	EXPANN_CUDA(h, cudaEventRecord(g.ev_q, st)); // the queries are in place / earlier users of d_keys_out are done
	const bool wait_merged = gr > 1 && g.merged_pending[par];
	// everything one kid needs for this call, issued on ITS stream; runs on the kid's own host thread (kid 0: this thread)
	auto kid_step = [&g, h, d_q, host_q, m, k, ef, gat, gr, d, par, wait_merged](size_t c, size_t q_lo, size_t mq) -> int {
		const size_t ir = c % gr;
		expann_gpu* kid = g.kids[c];
		DeviceGuard kg(kid->device);
#define MG_CUDA(call)                                                                            \
	do {                                                                                         \
		cudaError_t e_ = (call);                                                                 \
		if (e_ != cudaSuccess) { kid->err = std::string(#call) + ": " + cudaGetErrorString(e_); return EXPANN_ERR_CUDA; } \
	} while (0)
		const size_t cp = 2 * c + (size_t)par;
		cudaStream_t ks = par && g.kid_stream2[c] ? g.kid_stream2[c] : kid->stream;
		MG_CUDA(cudaStreamWaitEvent(ks, g.ev_q, 0));
		if (wait_merged) MG_CUDA(cudaStreamWaitEvent(ks, g.ev_merged[par], 0)); // last reader of this gather buffer
		const float* qk = d_q + q_lo * d;
		if (host_q) { // this kid's block of the queries comes straight from the caller's page-locked memory
			MG_CUDA(g.kid_q[cp].ensure(mq * d * sizeof(float)));
			MG_CUDA(cudaMemcpyAsync(g.kid_q[cp].p, host_q + q_lo * d, mq * d * sizeof(float), cudaMemcpyHostToDevice, ks));
			qk = g.kid_q[cp].as<float>();
		} else if (!g.same_dev[c]) { // this kid's block of the queries comes over NVLink from GPU 0
			MG_CUDA(g.kid_q[cp].ensure(mq * d * sizeof(float)));
			MG_CUDA(cudaMemcpyPeerAsync(g.kid_q[cp].p, kid->device, qk, h->device, mq * d * sizeof(float), ks));
			qk = g.kid_q[cp].as<float>();
		}
		uint64_t* out = gat + (ir * m + q_lo) * k;
		int rc;
		if (g.direct[c])
			rc = query_device_impl(kid, qk, mq, k, ef, out, ks); // the kid's last kernel stores into GPU 0's memory
		else {
			MG_CUDA(g.kid_keys[cp].ensure(mq * k * 8));
			rc = query_device_impl(kid, qk, mq, k, ef, g.kid_keys[cp].as<uint64_t>(), ks);
			if (!rc) MG_CUDA(cudaMemcpyPeerAsync(out, h->device, g.kid_keys[cp].p, kid->device, mq * k * 8, ks));
		}
		if (rc) return rc;
		MG_CUDA(cudaEventRecord(g.ev_done[cp], ks));
#undef MG_CUDA
		return EXPANN_OK;
	};
	std::vector<size_t> active;
	size_t q_lo0 = 0, mq0 = 0;
	bool run0 = false;
	for (size_t c = 0; c < g.kids.size(); ++c) {
		size_t q_lo, q_hi;
		mg_query_block(g, m, c / gr, &q_lo, &q_hi);
		if (q_lo >= q_hi) continue;
		const size_t mq = q_hi - q_lo;
		if (!g.same_dev[c]) g.peer_bytes += (host_q ? 0 : mq * d * sizeof(float)) + mq * k * 8;
		active.push_back(c);
		if (c == 0) { run0 = true; q_lo0 = q_lo; mq0 = mq; }
		else g.workers[c]->post([kid_step, c, q_lo, mq] { return kid_step(c, q_lo, mq); });
	}
	int rc_all = run0 ? kid_step(0, q_lo0, mq0) : EXPANN_OK;
	size_t bad = 0;
	for (size_t c : active) {
		if (c == 0) continue;
		const int rc = g.workers[c]->wait(); // every kid's launches (and its done event) are issued before GPU 0's stream waits on them
		if (rc && !rc_all) { rc_all = rc; bad = c; }
	}
	if (rc_all) return kid_fail(h, g.kids[bad], rc_all, bad);
	for (size_t c : active) EXPANN_CUDA(h, cudaStreamWaitEvent(st, g.ev_done[2 * c + (size_t)par], 0));
	if (gr > 1) {
		if (int rc = launch_merge(h, gat, m, gr, k, k, m * k, k, 0, d_keys_out, st)) return rc;
		EXPANN_CUDA(h, cudaEventRecord(g.ev_merged[par], st));
		g.merged_pending[par] = true;
	}
	return EXPANN_OK;
}

int mg_query_device(expann_gpu* h, const float* d_q, size_t m, size_t k, size_t ef, uint64_t* d_keys_out, cudaStream_t st) {
	return mg_query_impl(h, d_q, nullptr, m, k, ef, d_keys_out, st);
}

int mg_query(expann_gpu* h, const float* q, size_t m, size_t k, size_t ef, uint64_t* ids_out, float* dist_out, uint64_t* counts_out) {
	Multi& g = *h->mg;
	if (!h->built) return fail(h, EXPANN_ERR_STATE, "query before build()");
	if (k == 0) return fail(h, EXPANN_ERR_ARG, "k must be > 0");
	if (m == 0) return EXPANN_OK;
	DeviceGuard rg(h->device);
	cudaStream_t st = h->stream;
	const size_t qbytes = m * h->d * sizeof(float), nk = m * k;
	const size_t off_d = nk * 8, off_c = round_up_sz(off_d + nk * 4, 8), res_bytes = off_c + m * 8;
	EXPANN_CUDA(h, g.q_dev.ensure(qbytes));
	EXPANN_CUDA(h, g.keys.ensure(nk * 8));
	EXPANN_CUDA(h, g.res.ensure(res_bytes));
	uint8_t* d_res = g.res.as<uint8_t>();
	const bool q_pinned = is_pinned_host(q), ids_pinned = is_pinned_host(ids_out);
	const bool direct = ids_pinned && (!dist_out || is_pinned_host(dist_out)) && (!counts_out || is_pinned_host(counts_out));
	if (!q_pinned) {
		EXPANN_CUDA(h, h->h_q.ensure(qbytes));
		std::memcpy(h->h_q.p, q, qbytes);
	}
	const float* hq = q_pinned ? q : static_cast<const float*>(h->h_q.p);
	if (g.gq > 1) {
		// query groups: every GPU needs only its block -- copied from the host by each GPU itself, all PCIe links side by side
		if (int rc = mg_query_impl(h, nullptr, hq, m, k, ef, g.keys.as<uint64_t>(), st)) return rc;
	} else {
		// every GPU needs every query: ONE H2D copy to GPU 0, the others pull it over NVLink
		EXPANN_CUDA(h, cudaMemcpyAsync(g.q_dev.p, hq, qbytes, cudaMemcpyHostToDevice, st));
		if (int rc = mg_query_impl(h, g.q_dev.as<float>(), nullptr, m, k, ef, g.keys.as<uint64_t>(), st)) return rc;
	}
	if (int rc = launch_decode(h, g.keys.as<uint64_t>(), m, k, 0, reinterpret_cast<uint64_t*>(d_res), reinterpret_cast<float*>(d_res + off_d),
	                           reinterpret_cast<uint64_t*>(d_res + off_c), st))
		return rc;
	if (direct) {
		EXPANN_CUDA(h, cudaMemcpyAsync(ids_out, d_res, nk * 8, cudaMemcpyDeviceToHost, st));
		if (dist_out) EXPANN_CUDA(h, cudaMemcpyAsync(dist_out, d_res + off_d, nk * 4, cudaMemcpyDeviceToHost, st));
		if (counts_out) EXPANN_CUDA(h, cudaMemcpyAsync(counts_out, d_res + off_c, m * 8, cudaMemcpyDeviceToHost, st));
		EXPANN_CUDA(h, cudaStreamSynchronize(st));
	} else {
		const size_t need = counts_out ? res_bytes : (dist_out ? off_d + nk * 4 : nk * 8);
		EXPANN_CUDA(h, h->h_ids.ensure(res_bytes));
		EXPANN_CUDA(h, cudaMemcpyAsync(h->h_ids.p, d_res, need, cudaMemcpyDeviceToHost, st));
		EXPANN_CUDA(h, cudaStreamSynchronize(st));
		const uint8_t* r = static_cast<const uint8_t*>(h->h_ids.p);
		std::memcpy(ids_out, r, nk * 8);
		if (dist_out) std::memcpy(dist_out, r + off_d, nk * 4);
		if (counts_out) std::memcpy(counts_out, r + off_c, m * 8);
	}
	return EXPANN_OK;
}

int mg_set_option(expann_gpu* h, const char* name, int64_t value) {
	Multi& g = *h->mg;
	const std::string nm(name);
	if (nm == "mgpu_row_groups") {
		if (value < 0) return fail(h, EXPANN_ERR_ARG, "mgpu_row_groups must be >= 0 (0 = automatic)");
		if (g.materialized) return fail(h, EXPANN_ERR_STATE, "mgpu_row_groups must be set before the shards are created (build / build_graph / load_index)");
		g.opt_row_groups = (int)value;
		return EXPANN_OK;
	}
	for (size_t c = 0; c < g.kids.size(); ++c)
		if (int rc = expann_gpu_set_option(g.kids[c], name, value)) return kid_fail(h, g.kids[c], rc, c);
	for (auto& o : g.options)
		if (o.first == nm) { o.second = value; return EXPANN_OK; }
	g.options.emplace_back(nm, value);
	return EXPANN_OK;
}

int mg_set_timing(expann_gpu* h, int on) {
	Multi& g = *h->mg;
	g.timing = on != 0;
	for (expann_gpu* kid : g.kids) expann_gpu_set_timing(kid, on);
	return EXPANN_OK;
}

int mg_stats(expann_gpu* h, expann_gpu_stats_t* out) {
	Multi& g = *h->mg;
	expann_gpu_stats_t s = h->stats; // GPU 0's own merge / decode launches
	s.n = h->n;
	s.d = h->d;
	s.d_pad = h->d_pad;
	s.hot_time_ns = 0;
	s.hot_timed_launches = 0;
	for (size_t c = 0; c < g.kids.size(); ++c) {
		expann_gpu_stats_t ks;
		if (int rc = expann_gpu_stats(g.kids[c], &ks)) return kid_fail(h, g.kids[c], rc, c);
		s.scan_launches += ks.scan_launches;
		s.merge_launches += ks.merge_launches;
		s.graph_launches += ks.graph_launches;
		s.aux_launches += ks.aux_launches;
		s.tc_launches += ks.tc_launches;
		s.tc_fallbacks += ks.tc_fallbacks;
		s.distcomps += ks.distcomps;
		if (c == 0) { s.last_nsplit = ks.last_nsplit; s.last_qtile = ks.last_qtile; s.tc_mode = ks.tc_mode; }
		// the kids run side by side: the step's hot time is the slowest kid's
		if (ks.hot_time_ns > s.hot_time_ns) { s.hot_time_ns = ks.hot_time_ns; s.hot_timed_launches = ks.hot_timed_launches; }
	}
	s.peer_bytes = g.peer_bytes;
	s.query_groups = (uint64_t)g.gq;
	s.row_groups = (uint64_t)g.gr;
	*out = s;
	return EXPANN_OK;
}

} // namespace expann
