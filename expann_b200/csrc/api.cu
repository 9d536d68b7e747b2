// C ABI of libexpann_gpu.so (declared in include/expann_gpu.h). Host-side state machine only; the kernels
// live in scan_f32.cuh / merge_topk.cuh / graph_search.cuh / quant_scan.cuh.
#include "engine.h"
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <new>

namespace expann {

static thread_local std::string g_create_err;

int fail(expann_gpu* h, int code, const std::string& msg) {
	if (h) h->err = msg; else g_create_err = msg;
	return code;
}
int cuda_fail(expann_gpu* h, cudaError_t e, const char* what) {
	return fail(h, EXPANN_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to the (kernel, device) pair, not to a handle or a thread: every
// launch site goes through here, the value only ever grows, and the check + set is one critical section (a per-thread or
// per-call-site cache lets a later, smaller request lower the limit under a launch that still needs the larger one).
int ensure_dyn_smem(expann_gpu* h, const void* kernel, size_t bytes) {
	static std::mutex mu;
	static std::map<std::pair<const void*, int>, size_t> granted;
	std::lock_guard<std::mutex> lk(mu);
	size_t& cur = granted[{ kernel, h->device }];
	if (bytes > cur) {
		EXPANN_CUDA(h, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
		cur = bytes;
	}
	return EXPANN_OK;
}

static cudaEvent_t get_event(expann_gpu* h) {
	if (!h->event_pool.empty()) { cudaEvent_t e = h->event_pool.back(); h->event_pool.pop_back(); return e; }
	cudaEvent_t e = nullptr;
	cudaEventCreate(&e);
	return e;
}
cudaEvent_t timing_begin(expann_gpu* h, cudaStream_t st) {
	if (!h->timing) return nullptr;
	cudaEvent_t e = get_event(h);
	if (e) cudaEventRecord(e, st);
	return e;
}
void timing_end(expann_gpu* h, cudaEvent_t start, cudaStream_t st) {
	if (!start) return;
	cudaEvent_t e = get_event(h);
	if (e) cudaEventRecord(e, st);
	h->timed.emplace_back(start, e);
}
static void timing_collect(expann_gpu* h) {
	for (auto& pr : h->timed) {
		float ms = 0.f;
		if (pr.second && cudaEventSynchronize(pr.second) == cudaSuccess &&
		    cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) {
			h->stats.hot_time_ns += (uint64_t)((double)ms * 1e6);
			h->stats.hot_timed_launches++;
		}
		if (pr.first) h->event_pool.push_back(pr.first);
		if (pr.second) h->event_pool.push_back(pr.second);
	}
	h->timed.clear();
}

static size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// grow the HBM base buffer to hold `rows` rows (zero-filled so the d..d_pad padding columns stay 0)
static int reserve_rows(expann_gpu* h, size_t rows) {
	if (rows <= h->cap_rows) return EXPANN_OK;
	size_t want = rows;
	if (h->cap_rows) want = std::max(rows, h->cap_rows + h->cap_rows / 2);
	float* nb = nullptr;
	EXPANN_CUDA(h, cudaMalloc(&nb, want * h->d_pad * sizeof(float)));
	cudaError_t e = cudaMemsetAsync(nb, 0, want * h->d_pad * sizeof(float), h->stream);
	if (e == cudaSuccess && h->n)
		e = cudaMemcpyAsync(nb, h->base, h->n * h->d_pad * sizeof(float), cudaMemcpyDeviceToDevice, h->stream);
	if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
	if (e != cudaSuccess) { cudaFree(nb); return cuda_fail(h, e, "reserve_rows"); }
	if (h->base) cudaFree(h->base);
	h->base = nb;
	h->cap_rows = want;
	return EXPANN_OK;
}

static int upload_rows(expann_gpu* h, const float* rows, size_t n, cudaMemcpyKind kind, cudaStream_t st) {
	if (int rc = reserve_rows(h, h->n + n)) return rc;
	EXPANN_CUDA(h, cudaMemcpy2DAsync(h->base + h->n * h->d_pad, h->d_pad * sizeof(float), rows, h->d * sizeof(float),
	                                 h->d * sizeof(float), n, kind, st));
	EXPANN_CUDA(h, cudaStreamSynchronize(st));
	h->n += n;
	h->built = false;
	return EXPANN_OK;
}

static int flush_staging(expann_gpu* h) {
	if (!h->staged_rows) return EXPANN_OK;
	size_t rows = h->staged_rows;
	h->staged_rows = 0;
	int rc = upload_rows(h, h->staging.data(), rows, cudaMemcpyHostToDevice, h->stream);
	h->staging.clear();
	return rc;
}

static int check_dim(expann_gpu* h, size_t d) {
	if (d == 0) return fail(h, EXPANN_ERR_ARG, "d must be > 0");
	if (h->d == 0) {
		h->d = d;
		h->d_pad = round_up(d, 32);
	} else if (h->d != d)
		return fail(h, EXPANN_ERR_ARG, "dimension mismatch with previously stored vectors");
	return EXPANN_OK;
}

// pad device queries [m][d] -> [m][d_pad] when d is not a multiple of 32 floats
static int padded_queries(expann_gpu* h, const float* d_q, size_t m, cudaStream_t st, const float** out) {
	if (h->d == h->d_pad) { *out = d_q; return EXPANN_OK; }
	EXPANN_CUDA(h, h->qpad.ensure(m * h->d_pad * sizeof(float)));
	EXPANN_CUDA(h, cudaMemsetAsync(h->qpad.p, 0, m * h->d_pad * sizeof(float), st));
	EXPANN_CUDA(h, cudaMemcpy2DAsync(h->qpad.p, h->d_pad * sizeof(float), d_q, h->d * sizeof(float), h->d * sizeof(float),
	                                 m, cudaMemcpyDefault, st)); // d_q may be mapped host memory (small calls)
	h->stats.aux_launches += 2;
	*out = h->qpad.as<float>();
	return EXPANN_OK;
}

static void swap_scratch(expann_gpu* h) {
	expann_gpu::Scratch& s = h->spare;
	std::swap(h->qpad, s.qpad);
	std::swap(h->partial, s.partial);
	std::swap(h->scan_cnt, s.scan_cnt);
	std::swap(h->tc_qstats, s.tc_qstats);
	std::swap(h->tc_qb, s.tc_qb);
	std::swap(h->tc_qn, s.tc_qn);
	std::swap(h->tc_thr, s.tc_thr);
	std::swap(h->tc_cnt, s.tc_cnt);
	std::swap(h->tc_cand, s.tc_cand);
	std::swap(h->tc_gthr, s.tc_gthr);
	std::swap(h->tc_sel, s.tc_sel);
	std::swap(h->tc_seed, s.tc_seed);
	h->set_active ^= 1;
}

// Which set of work buffers does a call on stream `st` use? The set this stream used last (stream order protects it);
// otherwise the set used longest ago, behind the event of its last call. The brute-force engine has two sets, so two
// streams never wait for one another; the other engines have one (a call on another stream waits for the last call).
static int scratch_for_stream(expann_gpu* h, cudaStream_t st) {
	const bool two = h->cfg.engine == EXPANN_ENGINE_BRUTE_FORCE;
	int a = h->set_active;
	if (h->set_used[a] && h->set_stream[a] == st) return EXPANN_OK;
	if (two) {
		const int o = a ^ 1;
		swap_scratch(h); // the other set: this stream's own from an earlier call, or the one used longest ago
		a = o;
		if (h->set_used[a] && h->set_stream[a] == st) return EXPANN_OK;
	}
	if (h->set_used[a]) EXPANN_CUDA(h, cudaStreamWaitEvent(st, h->set_done[a], 0));
	h->set_stream[a] = st;
	return EXPANN_OK;
}

int query_device_impl(expann_gpu* h, const float* d_q, size_t m, size_t k, size_t ef, uint64_t* d_keys_out,
                      cudaStream_t st) {
	if (!h->built) return fail(h, EXPANN_ERR_STATE, "query before build()");
	if (k == 0) return fail(h, EXPANN_ERR_ARG, "k must be > 0");
	if (m == 0) return EXPANN_OK;
	if (int rc = scratch_for_stream(h, st)) return rc;
	const float* qp = nullptr;
	int rc = padded_queries(h, d_q, m, st, &qp);
	if (!rc) switch (h->cfg.engine) {
		case EXPANN_ENGINE_BRUTE_FORCE:
			rc = launch_scan_f32(h, qp, m, k, d_keys_out, st);
			break;
		case EXPANN_ENGINE_GRAPH:
			rc = launch_graph_search(h, qp, m, k, ef, d_keys_out, st);
			break;
		case EXPANN_ENGINE_QUANT_SCAN:
			rc = launch_quant_scan(h, qp, m, k, d_keys_out, st);
			break;
		default:
			rc = fail(h, EXPANN_ERR_UNSUPPORTED, "engine kind not implemented");
		}
	// whatever was queued (even by a call that failed half way) is the set's last user
	const int a = h->set_active;
	if (!h->set_done[a] && cudaEventCreateWithFlags(&h->set_done[a], cudaEventDisableTiming) != cudaSuccess) {
		h->set_done[a] = nullptr;
		return rc ? rc : cuda_fail(h, cudaGetLastError(), "cudaEventCreate");
	}
	const cudaError_t e = cudaEventRecord(h->set_done[a], st);
	h->set_used[a] = true;
	if (e != cudaSuccess && !rc) return cuda_fail(h, e, "cudaEventRecord");
	return rc;
}

bool is_pinned_host(const void* p) {
	cudaPointerAttributes a{};
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
		cudaGetLastError(); // unregistered host memory reports an error on old drivers: not sticky, clear it
		return false;
	}
	return a.type == cudaMemoryTypeHost;
}

} // namespace expann

using namespace expann;

extern "C" {

const char* expann_gpu_version(void) { return "expann_b200 0.1; sm_100a; kernels: scan_topk_f32 merge_topk decode_keys"; }

int expann_gpu_create(const expann_gpu_cfg* cfg, expann_gpu_t** out) {
	if (!cfg || !out) return fail(nullptr, EXPANN_ERR_ARG, "null argument");
	*out = nullptr;
	if (cfg->engine != EXPANN_ENGINE_BRUTE_FORCE && cfg->engine != EXPANN_ENGINE_GRAPH &&
	    cfg->engine != EXPANN_ENGINE_QUANT_SCAN)
		return fail(nullptr, EXPANN_ERR_UNSUPPORTED, "engine kind not implemented");
	if (cfg->quantizer != 0 && cfg->quantizer != 1) return fail(nullptr, EXPANN_ERR_ARG, "bad quantizer");
	if (cfg->metric != EXPANN_METRIC_L2 && cfg->metric != EXPANN_METRIC_IP) return fail(nullptr, EXPANN_ERR_ARG, "bad metric");
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0)
		return fail(nullptr, EXPANN_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) +
		                                          " (this library has no CPU fallback)");
	int dev = cfg->device;
	if (dev < 0) cudaGetDevice(&dev);
	if (dev >= ndev) return fail(nullptr, EXPANN_ERR_ARG, "device ordinal out of range");
	if (cfg->n_devices > 1) { // GPUs dev .. dev + n_devices - 1 behind one handle (multi.cu)
		if (cfg->n_devices > 64) return fail(nullptr, EXPANN_ERR_ARG, "n_devices must be <= 64");
		int32_t list[64];
		for (int i = 0; i < cfg->n_devices; ++i) list[i] = dev + i;
		return mg_create(cfg, list, (size_t)cfg->n_devices, out);
	}
	cudaDeviceProp prop;
	e = cudaGetDeviceProperties(&prop, dev);
	if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceProperties");
	if (prop.major != 10) return fail(nullptr, EXPANN_ERR_UNSUPPORTED, "device is not sm_100 (kernels are built for sm_100a only)");
	expann_gpu* h = new (std::nothrow) expann_gpu();
	if (!h) return fail(nullptr, EXPANN_ERR_ARG, "out of host memory");
	h->cfg = *cfg;
	h->quantizer = cfg->quantizer;
	h->device = dev;
	DeviceGuard g(dev);
	e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
	if (e != cudaSuccess) { delete h; return cuda_fail(nullptr, e, "cudaStreamCreate"); }
	*out = h;
	return EXPANN_OK;
}

int expann_gpu_create_multi(const expann_gpu_cfg* cfg, const int32_t* devices, size_t n_devices, expann_gpu_t** out) {
	if (!cfg || !out) return fail(nullptr, EXPANN_ERR_ARG, "null argument");
	*out = nullptr;
	if (cfg->engine != EXPANN_ENGINE_BRUTE_FORCE && cfg->engine != EXPANN_ENGINE_GRAPH && cfg->engine != EXPANN_ENGINE_QUANT_SCAN)
		return fail(nullptr, EXPANN_ERR_UNSUPPORTED, "engine kind not implemented");
	if (cfg->quantizer != 0 && cfg->quantizer != 1) return fail(nullptr, EXPANN_ERR_ARG, "bad quantizer");
	if (cfg->metric != EXPANN_METRIC_L2 && cfg->metric != EXPANN_METRIC_IP) return fail(nullptr, EXPANN_ERR_ARG, "bad metric");
	return mg_create(cfg, devices, n_devices, out);
}

void expann_gpu_destroy(expann_gpu_t* h) {
	if (!h) return;
	if (h->mg) mg_destroy(h);
	DeviceGuard g(h->device);
	if (h->stream) { cudaStreamSynchronize(h->stream); cudaStreamDestroy(h->stream); }
	timing_collect(h);
	for (cudaEvent_t e : h->event_pool) cudaEventDestroy(e);
	if (h->tc_flag_event) cudaEventDestroy(h->tc_flag_event);
	h->tc_flag_host.release();
	if (h->base) cudaFree(h->base);
	for (DevBuf* b : { &h->tc_xb, &h->tc_xn, &h->tc_stats, &h->tc_qstats, &h->tc_qb, &h->tc_qn, &h->tc_thr, &h->tc_cnt, &h->tc_cand, &h->tc_gthr, &h->tc_sel, &h->tc_xx, &h->tc_xx32, &h->tc_seed, &h->tc_fbcount, &h->scan_cnt })
		b->release();
	for (DevBuf* b : { &h->q_codes, &h->q_cnorm, &h->q_mm, &h->qq_codes, &h->qq_norm, &h->q_cand, &h->q_cnb, &h->q_thrfin, &h->q_zero }) b->release();
	for (DevBuf* b : { &h->g_bottom, &h->g_upper, &h->g_upper_idx, &h->g_upper_nl, &h->g_visited, &h->g_epoch, &h->g_codes, &h->g_up_meta,
	                   &h->g_upper_c, &h->g_up_vec, &h->g_blk_bottom, &h->g_blk_upper })
		b->release();
	for (DevBuf* b : { &h->qpad, &h->partial, &h->keys, &h->ids, &h->dists, &h->counts, &h->h_q, &h->h_ids, &h->h_dists,
	                   &h->h_counts })
		b->release();
	for (DevBuf* b : { &h->spare.qpad, &h->spare.partial, &h->spare.scan_cnt, &h->spare.tc_qstats, &h->spare.tc_qb, &h->spare.tc_qn, &h->spare.tc_thr,
	                   &h->spare.tc_cnt, &h->spare.tc_cand, &h->spare.tc_gthr, &h->spare.tc_sel, &h->spare.tc_seed })
		b->release();
	for (cudaEvent_t e : h->set_done)
		if (e) cudaEventDestroy(e);
	delete h;
}

const char* expann_gpu_last_error(const expann_gpu_t* h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int expann_gpu_add(expann_gpu_t* h, const float* rows, size_t n, size_t d) {
	if (!h || (!rows && n)) return fail(h, EXPANN_ERR_ARG, "null argument");
	if (h->mg) return mg_add(h, rows, n, d, false, nullptr);
	if (int rc = check_dim(h, d)) return rc;
	if (n == 0) return EXPANN_OK;
	DeviceGuard g(h->device);
	const size_t small = 4096; // rows; per-vector store_vector calls are batched on the host
	if (n < small) {
		h->staging.insert(h->staging.end(), rows, rows + n * d);
		h->staged_rows += n;
		h->built = false;
		if (h->staging.size() * sizeof(float) >= (size_t)64 << 20) return flush_staging(h);
		return EXPANN_OK;
	}
	if (int rc = flush_staging(h)) return rc;
	return upload_rows(h, rows, n, cudaMemcpyHostToDevice, h->stream);
}

int expann_gpu_add_device(expann_gpu_t* h, const float* d_rows, size_t n, size_t d, void* stream) {
	if (!h || (!d_rows && n)) return fail(h, EXPANN_ERR_ARG, "null argument");
	if (h->mg) return mg_add(h, d_rows, n, d, true, (cudaStream_t)stream);
	if (int rc = check_dim(h, d)) return rc;
	if (n == 0) return EXPANN_OK;
	DeviceGuard g(h->device);
	if (int rc = flush_staging(h)) return rc;
	cudaStream_t st = (cudaStream_t)stream;
	EXPANN_CUDA(h, cudaStreamSynchronize(st));
	return upload_rows(h, d_rows, n, cudaMemcpyDeviceToDevice, h->stream);
}

int expann_gpu_build(expann_gpu_t* h) {
	if (!h) return EXPANN_ERR_ARG;
	if (h->mg) return mg_build(h);
	DeviceGuard g(h->device);
	if (int rc = flush_staging(h)) return rc;
	if (h->n == 0) return fail(h, EXPANN_ERR_STATE, "build() on an empty engine (the reference asserts n > 0)");
	if (h->n + h->cfg.id_offset > 0xFFFFFFFEull) return fail(h, EXPANN_ERR_UNSUPPORTED, "ids must fit 32 bits on device");
	if (int rc = make_base_tensor_map(h)) return rc;
	if (h->cfg.engine == EXPANN_ENGINE_GRAPH)
		if (int rc = graph_finalize(h)) return rc;
	if (h->cfg.engine == EXPANN_ENGINE_QUANT_SCAN)
		if (int rc = quant_finalize(h)) return rc;
	if (int rc = tc_finalize(h)) return rc;
	if (h->cfg.engine == EXPANN_ENGINE_BRUTE_FORCE) scan_noalloc_setup(h); // address of the static scratch array on this device
	h->stats.n = h->n;
	h->stats.tc_mode = h->tc_mode;
	h->stats.d = h->d;
	h->stats.d_pad = h->d_pad;
	h->built = true;
	return EXPANN_OK;
}

int expann_gpu_query_device(expann_gpu_t* h, const float* d_q, size_t m, size_t k, size_t ef, uint64_t* d_keys_out,
                            void* stream) {
	if (!h || (m && (!d_q || !d_keys_out))) return fail(h, EXPANN_ERR_ARG, "null argument");
	if (h->mg) return mg_query_device(h, d_q, m, k, ef, d_keys_out, (cudaStream_t)stream);
	DeviceGuard g(h->device);
	return query_device_impl(h, d_q, m, k, ef, d_keys_out, (cudaStream_t)stream);
}

int expann_gpu_scan_device_noalloc(expann_gpu_t* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, void* stream) {
	if (!h || (m && (!d_q || !d_keys_out))) return fail(h, EXPANN_ERR_ARG, "null argument");
	if (h->mg) return fail(h, EXPANN_ERR_UNSUPPORTED, "scan_device_noalloc: single-device handles only");
	DeviceGuard g(h->device);
	return scan_noalloc(h, d_q, m, k, d_keys_out, (cudaStream_t)stream);
}

constexpr size_t SMALL_CALL_BYTES = 16 << 10; // queries / results of a call up to this size skip the copy engines

int expann_gpu_query(expann_gpu_t* h, const float* q, size_t m, size_t k, size_t ef, uint64_t* ids_out, float* dist_out,
                     uint64_t* counts_out) {
	if (!h || (m && (!q || !ids_out))) return fail(h, EXPANN_ERR_ARG, "null argument");
	if (h->mg) return mg_query(h, q, m, k, ef, ids_out, dist_out, counts_out);
	if (!h->built) return fail(h, EXPANN_ERR_STATE, "query before build()");
	if (k == 0) return fail(h, EXPANN_ERR_ARG, "k must be > 0");
	if (m == 0) return EXPANN_OK;
	DeviceGuard g(h->device);
	cudaStream_t st = h->stream;
	const size_t qbytes = m * h->d * sizeof(float), nk = m * k;
	// results leave the device as ONE block [ids nk x 8 | keys nk x 4 | counts m x 8] (8-byte aligned parts): a staged
	// (pageable) caller gets one device-to-host copy instead of three
	const size_t off_d = nk * 8, off_c = round_up(off_d + nk * 4, 8), res_bytes = off_c + m * 8;
	EXPANN_CUDA(h, h->h_q.ensure(qbytes));
	EXPANN_CUDA(h, h->h_ids.ensure(res_bytes));
	EXPANN_CUDA(h, h->keys.ensure(nk * 8 + qbytes + 256)); // keys | raw queries
	EXPANN_CUDA(h, h->ids.ensure(res_bytes));
	uint8_t* d_res = h->ids.as<uint8_t>();
	uint64_t* d_ids = reinterpret_cast<uint64_t*>(d_res);
	float* d_dist = reinterpret_cast<float*>(d_res + off_d);
	uint64_t* d_cnt = reinterpret_cast<uint64_t*>(d_res + off_c);
	// page-locked caller buffers (cudaHostAlloc / cudaHostRegister / torch pin_memory) are used by the copy engines
	// directly; pageable ones go through the handle's pinned staging buffers
	const bool q_pinned = is_pinned_host(q), ids_pinned = is_pinned_host(ids_out);
	const bool dist_pinned = dist_out && is_pinned_host(dist_out), counts_pinned = counts_out && is_pinned_host(counts_out);
	const bool direct = ids_pinned && (!dist_out || dist_pinned) && (!counts_out || counts_pinned);
	if (qbytes <= SMALL_CALL_BYTES && res_bytes <= SMALL_CALL_BYTES) {
		// few queries per call (the reference harness's own loop is one per call, src/basic_bench.h:82-126): no copy
		// engine behind the kernels — the decode kernel writes the results into the handle's page-locked staging buffer
		// (mapped into the device's address space), and the graph search reads its queries from one; up to four copy
		// operations and their latencies are replaced by small host memcpys
		const bool q_pin = is_pinned_host(q);
		if (!q_pin) std::memcpy(h->h_q.p, q, qbytes);
		const float* qsrc = static_cast<const float*>(h->h_q.p);
		if (h->cfg.engine != EXPANN_ENGINE_GRAPH) {
			// the scans read a query once per CTA: those reads stay in HBM (the graph search reads it exactly once)
			float* d_q = reinterpret_cast<float*>(h->keys.as<uint8_t>() + round_up(nk * 8, 256));
			EXPANN_CUDA(h, cudaMemcpyAsync(d_q, q_pin ? (const void*)q : h->h_q.p, qbytes, cudaMemcpyHostToDevice, st));
			qsrc = d_q;
		} else if (q_pin)
			std::memcpy(h->h_q.p, q, qbytes);
		uint8_t* r = static_cast<uint8_t*>(h->h_ids.p);
		h->fused = false;
		if (h->cfg.engine == EXPANN_ENGINE_GRAPH) {
			h->fuse_ids = reinterpret_cast<uint64_t*>(r);
			h->fuse_dists = reinterpret_cast<float*>(r + off_d);
			h->fuse_counts = reinterpret_cast<uint64_t*>(r + off_c);
		}
		const int qrc = query_device_impl(h, qsrc, m, k, ef, h->keys.as<uint64_t>(), st);
		h->fuse_ids = nullptr;
		if (qrc) return qrc;
		if (!h->fused)
			if (int rc = launch_decode(h, h->keys.as<uint64_t>(), m, k, 0, reinterpret_cast<uint64_t*>(r), reinterpret_cast<float*>(r + off_d),
			                           reinterpret_cast<uint64_t*>(r + off_c), st))
				return rc;
		EXPANN_CUDA(h, cudaStreamSynchronize(st));
		std::memcpy(ids_out, r, nk * 8);
		if (dist_out) std::memcpy(dist_out, r + off_d, nk * 4);
		if (counts_out) std::memcpy(counts_out, r + off_c, m * 8);
		return EXPANN_OK;
	}
	if (!q_pinned) std::memcpy(h->h_q.p, q, qbytes);
	float* d_q = reinterpret_cast<float*>(h->keys.as<uint8_t>() + round_up(nk * 8, 256));
	EXPANN_CUDA(h, cudaMemcpyAsync(d_q, q_pinned ? (const void*)q : h->h_q.p, qbytes, cudaMemcpyHostToDevice, st));
	if (int rc = query_device_impl(h, d_q, m, k, ef, h->keys.as<uint64_t>(), st)) return rc;
	if (int rc = launch_decode(h, h->keys.as<uint64_t>(), m, k, 0, d_ids, d_dist, d_cnt, st)) return rc;
	if (direct) {
		EXPANN_CUDA(h, cudaMemcpyAsync(ids_out, d_ids, nk * 8, cudaMemcpyDeviceToHost, st));
		if (dist_out) EXPANN_CUDA(h, cudaMemcpyAsync(dist_out, d_dist, nk * 4, cudaMemcpyDeviceToHost, st));
		if (counts_out) EXPANN_CUDA(h, cudaMemcpyAsync(counts_out, d_cnt, m * 8, cudaMemcpyDeviceToHost, st));
		EXPANN_CUDA(h, cudaStreamSynchronize(st));
	} else {
		const size_t need = counts_out ? res_bytes : (dist_out ? off_d + nk * 4 : nk * 8);
		EXPANN_CUDA(h, cudaMemcpyAsync(h->h_ids.p, d_res, need, cudaMemcpyDeviceToHost, st));
		EXPANN_CUDA(h, cudaStreamSynchronize(st));
		const uint8_t* r = static_cast<const uint8_t*>(h->h_ids.p);
		std::memcpy(ids_out, r, nk * 8);
		if (dist_out) std::memcpy(dist_out, r + off_d, nk * 4);
		if (counts_out) std::memcpy(counts_out, r + off_c, m * 8);
	}
	return EXPANN_OK;
}

int expann_gpu_merge_device(expann_gpu_t* h, const uint64_t* d_keys_in, size_t m, size_t nlists, size_t klist,
                            size_t stride_q, size_t stride_s, size_t k, uint64_t* d_keys_out, void* stream) {
	if (!h || (m && (!d_keys_in || !d_keys_out))) return fail(h, EXPANN_ERR_ARG, "null argument");
	DeviceGuard g(h->device);
	return launch_merge(h, d_keys_in, m, nlists, klist, stride_q, stride_s, k, 0, d_keys_out, (cudaStream_t)stream);
}

int expann_gpu_decode_device(expann_gpu_t* h, const uint64_t* d_keys, size_t m, size_t k, uint64_t* d_ids, float* d_dist,
                             uint64_t* d_counts, void* stream) {
	if (!h || (m && (!d_keys || !d_ids))) return fail(h, EXPANN_ERR_ARG, "null argument");
	DeviceGuard g(h->device);
	return launch_decode(h, d_keys, m, k, 0, d_ids, d_dist, d_counts, (cudaStream_t)stream);
}

int expann_gpu_load_index(expann_gpu_t* h, const char* path) {
	if (!h || !path) return fail(h, EXPANN_ERR_ARG, "null argument");
	if (h->cfg.engine != EXPANN_ENGINE_GRAPH) return fail(h, EXPANN_ERR_STATE, "load_index: not a graph engine");
	if (h->mg) return mg_load_index(h, path);
	DeviceGuard g(h->device);
	std::vector<float> vectors;
	size_t n = 0, d = 0;
	if (int rc = graph_load_index(h, path, vectors, n, d)) return rc;
	// like read_index in the reference (src/antitopo_engine.h:312-313,476-481): the file's vectors replace any stored ones
	// The base buffer is sized in rows of the OLD d_pad and its padding columns are only zero for the old d: drop it, so
	// that reserve_rows() allocates and zero-fills for the file's dimension (a smaller d would overflow the buffer, the
	// same d_pad with a smaller d would leave stale values in the padding the kernels rely on being zero).
	h->staging.clear();
	h->staged_rows = 0;
	EXPANN_CUDA(h, cudaStreamSynchronize(h->stream));
	if (h->base) { cudaFree(h->base); h->base = nullptr; }
	h->cap_rows = 0;
	h->n = 0;
	h->d = 0;
	h->built = false;
	if (int rc = check_dim(h, d)) return rc;
	return upload_rows(h, vectors.data(), n, cudaMemcpyHostToDevice, h->stream);
}

int expann_gpu_build_graph(expann_gpu_t* h, uint64_t M, uint64_t M0, uint64_t K, float ortho_factor, float ortho_bias,
                           uint64_t prune_overflow, uint64_t seed) {
	if (!h) return EXPANN_ERR_ARG;
	if (h->cfg.engine != EXPANN_ENGINE_GRAPH) return fail(h, EXPANN_ERR_STATE, "build_graph: not a graph engine");
	if (h->mg) return mg_build_graph(h, M, M0, K, ortho_factor, ortho_bias, prune_overflow, seed);
	DeviceGuard g(h->device);
	if (int rc = flush_staging(h)) return rc;
	return graph_build_gpu(h, M, M0, K, ortho_factor, ortho_bias, prune_overflow, seed);
}

int expann_gpu_save_index(expann_gpu_t* h, const char* path) {
	if (!h || !path) return EXPANN_ERR_ARG;
	if (h->cfg.engine != EXPANN_ENGINE_GRAPH) return fail(h, EXPANN_ERR_STATE, "save_index: not a graph engine");
	if (h->mg) return fail(h, EXPANN_ERR_UNSUPPORTED, "save_index: a multi-device handle holds one sub-graph per GPU (save from single-device handles)");
	DeviceGuard g(h->device);
	if (int rc = flush_staging(h)) return rc;
	return graph_save_index(h, path);
}

int expann_gpu_set_graph(expann_gpu_t* h, size_t n, const uint32_t* nlayers, const uint64_t* offsets, const uint32_t* edges,
                         uint64_t starting_vertex, uint64_t max_layer) {
	if (!h || !nlayers || !offsets || !edges) return fail(h, EXPANN_ERR_ARG, "null argument");
	if (h->cfg.engine != EXPANN_ENGINE_GRAPH) return fail(h, EXPANN_ERR_STATE, "set_graph: not a graph engine");
	if (h->mg) return fail(h, EXPANN_ERR_UNSUPPORTED, "set_graph: multi-device handles take the graph from load_index (replicas) or build_graph (one sub-graph per shard)");
	return graph_set_csr(h, n, nlayers, offsets, edges, starting_vertex, max_layer);
}

int expann_gpu_get_quantizer(const expann_gpu_t* h, float* scale, float* offset) {
	if (!h || !scale || !offset) return EXPANN_ERR_ARG;
	*scale = h->q_scale;
	*offset = h->q_offset;
	return EXPANN_OK;
}

int expann_gpu_set_timing(expann_gpu_t* h, int on) {
	if (!h) return EXPANN_ERR_ARG;
	if (h->mg) return mg_set_timing(h, on);
	DeviceGuard g(h->device);
	timing_collect(h);
	h->timing = on != 0;
	return EXPANN_OK;
}

int expann_gpu_set_option(expann_gpu_t* h, const char* name, int64_t value) {
	if (!h || !name) return EXPANN_ERR_ARG;
	const std::string nm(name);
	if (h->mg && nm == "mgpu_row_groups") return mg_set_option(h, name, value);
	if (nm == "qtile") {
		if (value != 0 && value != 1 && value != 4 && value != 8) return fail(h, EXPANN_ERR_ARG, "qtile must be 0, 1, 4 or 8");
		h->cfg.qtile = (int32_t)value;
	} else if (nm == "kprime") {
		if (value < 0 || value > 128) return fail(h, EXPANN_ERR_ARG, "kprime must be in 0..128");
		h->opt_kprime = (int)value;
	} else if (nm == "tc_pair") {
		h->opt_tc_pair = value != 0;
	} else if (nm == "tc_astream") {
		if (value < 0 || value > 2) return fail(h, EXPANN_ERR_ARG, "tc_astream must be 0, 1 or 2");
		h->opt_tc_astream = (int)value;
	} else if (nm == "graph_occ") {
		if (value != 2 && value != 4 && value != 5 && value != 8) return fail(h, EXPANN_ERR_ARG, "graph_occ must be 2, 4, 5 or 8");
		if (h->built) return fail(h, EXPANN_ERR_STATE, "graph_occ must be set before build()");
		h->opt_graph_occ = (int)value;
	} else if (nm == "graph_cta_batch") {
		if (value < -1) return fail(h, EXPANN_ERR_ARG, "graph_cta_batch must be >= -1 (-1 = automatic)");
		h->opt_graph_cta_batch = (int)value;
	} else if (nm == "graph_cta_width") {
		if (value != 1 && value != 2) return fail(h, EXPANN_ERR_ARG, "graph_cta_width must be 1 or 2");
		h->opt_graph_cta_width = (int)value;
	} else if (nm == "graph_hash") {
		h->opt_graph_hash = value != 0;
	} else if (nm == "graph_inline") {
		if (value < -1 || value > 2) return fail(h, EXPANN_ERR_ARG, "graph_inline must be -1 (automatic), 0, 1 or 2");
		h->opt_graph_inline = (int)value; // before build(): whether the blocks are built; after it: whether blocks that exist are used
	} else if (nm == "graph_inline_batch") {
		h->opt_graph_inline_batch = value != 0;
	} else if (nm == "graph_upper_compact") {
		h->opt_graph_upper_compact = value != 0;
	} else if (nm == "tc_boot") {
		h->opt_tc_boot = (int)value; // 0 off, 1 every format but the split (default), 2 all formats
	} else if (nm == "tc_boot_tiles") {
		if (value < 1 || value > 4096) return fail(h, EXPANN_ERR_ARG, "tc_boot_tiles must be 1..4096");
		h->opt_tc_boot_tiles = (int)value;
	} else if (nm == "tc_real") {
		if (value != 0 && value != 1) return fail(h, EXPANN_ERR_ARG, "tc_real must be 0 or 1");
		if (h->built) return fail(h, EXPANN_ERR_STATE, "tc_real must be set before build()");
		h->opt_tc_real = (int)value;
	} else if (nm == "quant_tc") {
		h->opt_quant_tc = value != 0;
	} else if (nm == "quant_tc_min_batch") {
		if (value < 1) return fail(h, EXPANN_ERR_ARG, "quant_tc_min_batch must be >= 1");
		h->opt_quant_tc_min_batch = (int)value;
	} else if (nm == "tc_tf32") {
		if (h->built) return fail(h, EXPANN_ERR_STATE, "tc_tf32 must be set before build()");
		h->opt_tc_tf32 = value != 0;
	} else if (nm == "tc_v1") {
		h->opt_tc_v1 = value != 0;
	} else if (nm == "tc") {
		h->opt_tc = value != 0;
	} else if (nm == "tc_min_batch") {
		if (value < 1) return fail(h, EXPANN_ERR_ARG, "tc_min_batch must be >= 1");
		h->opt_tc_min_batch = (int)value;
	} else if (nm == "nsplit") {
		if (value < 0) return fail(h, EXPANN_ERR_ARG, "nsplit must be >= 0");
		h->opt_nsplit = (int)value;
	} else
		return fail(h, EXPANN_ERR_ARG, "unknown option: " + nm);
	if (h->mg) return mg_set_option(h, name, value); // validated above on the parent's own fields; now every GPU's handle
	return EXPANN_OK;
}

int expann_gpu_stats(expann_gpu_t* h, expann_gpu_stats_t* out) {
	if (!h || !out) return EXPANN_ERR_ARG;
	if (h->mg) return mg_stats(h, out);
	DeviceGuard g(h->device);
	timing_collect(h);
	if (h->cfg.engine == EXPANN_ENGINE_GRAPH) graph_read_distcomps(h);
	if (h->tc_fbcount.cap >= 4) { // redo launches of K3's second form are counted on the device
		uint32_t c = 0;
		if (cudaDeviceSynchronize() == cudaSuccess && cudaMemcpy(&c, h->tc_fbcount.p, 4, cudaMemcpyDeviceToHost) == cudaSuccess)
			h->stats.tc_fallbacks = c;
	}
	*out = h->stats;
	return EXPANN_OK;
}

} // extern "C"
