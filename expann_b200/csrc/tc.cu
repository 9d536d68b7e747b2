// Host side of K3: operand preparation at build(), filter launches, selection, fp32 re-rank.
//
// Forms (all share the operand formats, the bound argument below and the re-rank):
//   * second form, k <= 16 (tc_batch2): ONE filter launch with in-kernel per-query bounds shared between row splits --
//       pair form  (tc_filter3.cuh, clusters of two CTAs, cta_group::2): batches > 128 queries; query tile resident in
//                  shared memory, or streamed through the ring beside the row tiles when the rows are too long for that
//                  (d > 128 real-valued / d > 384 bf16-exact: GIST's 960); optional seeding launch for big batches;
//       single-CTA form (tc_filter2.cuh): batches <= 128 queries;
//     then tc_select2 (exact k-th smallest a' per query, survivors) and K4 (rerank.cuh). A batch whose flag is up
//     (buffer overflow, queries not bf16-exact, too many survivors) is redone by the SIMT scan, queued behind the flag
//     on the device (ScanParams::gate) -- no host round trip.
//   * first form, 16 < k <= 128 (tc_batch, tc_filter.cuh): row chunks of doubling size with the bound refreshed between
//     launches; needs a resident query tile and >= 512 queries. The graph build's bottom-layer kNN (k = K + 1) uses it.
//
// Why the result is exact. Let a'(x) be the tensor-core value and e >= |a'(x) - a'_true(x)| the error bound of the
// operand format (0 for bf16-exact integer data; eps_rel*|q|*max|x| for the two-term split). After any prefix of
// the rows, let A_k be the k-th smallest a' seen. k seen rows have a'_true <= A_k + e, hence the final k-th true
// value D_k <= A_k + e, and every row of the true top-k has a' <= D_k + e <= A_k + 2e =: thr. Rows are processed
// in chunks of doubling size; each chunk is filtered with the thr of the rows before it (first chunk: +inf) and
// thr only decreases, so the buffers always hold the true top-k. K4 then recomputes the survivors with the same
// fp32 diff-form arithmetic as the SIMT scan and orders them by (key, id). A buffer overflow (adversarial row order)
// is detected and the batch is redone by the SIMT scan.
#include "engine.h"
#include "rerank.cuh"
#include "tc_filter.cuh"
#include "tc_filter2.cuh"
#include "tc_filter3.cuh"
#include "tc_prep.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>

namespace expann {

int encode_tensor_map_bf16(expann_gpu* h, CUtensorMap* out, void* base, uint64_t inner, uint64_t rows, uint32_t box_rows,
                           uint32_t box_inner = 64);
int launch_scan_f32_simt(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st,
                         const uint32_t* gate = nullptr, uint32_t* gate_count = nullptr);

constexpr uint32_t TC_CAP = 4096, TC_FIRST_ROWS = 2048, TC_PAD = 256, TC_QBATCH = 16384;
constexpr int TC_NSTAGE_EXACT = 4, TC_NSTAGE_SPLIT = 3;
// second form (tc_filter2.cuh): per-(query, split) candidate buffers of TC2_CAP keys, TC_PAD survivors per query
constexpr uint32_t TC2_CAP = 512;
constexpr int TC2_NSTAGE = 6;

static size_t tc_smem_bytes(uint32_t kd, int nstage) {
	return 1024 + (size_t)(kd / TC_BK) * TC_A_KB_BYTES + (size_t)nstage * TC_B_STAGE_BYTES + (1 + 2 * nstage + 4) * 8 + 16;
}

static float __uint_as_float_host(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

static int read_stats(expann_gpu* h, DevBuf& buf, TcStats* out, cudaStream_t st) {
	EXPANN_CUDA(h, cudaMemcpyAsync(out, buf.p, sizeof(TcStats), cudaMemcpyDeviceToHost, st));
	EXPANN_CUDA(h, cudaStreamSynchronize(st));
	return EXPANN_OK;
}
static int reset_stats(expann_gpu* h, DevBuf& buf, cudaStream_t st) {
	EXPANN_CUDA(h, buf.ensure(sizeof(TcStats)));
	const TcStats init = { f32_orderable(3.4e38f), f32_orderable(-3.4e38f), 0u, f32_orderable(0.f) };
	EXPANN_CUDA(h, cudaMemcpyAsync(buf.p, &init, sizeof(init), cudaMemcpyHostToDevice, st));
	return EXPANN_OK;
}

// operand row length (bf16 elements) of a format: one term per value in the exact (1) and one-term (3) formats,
// [xh | xl | xh] in the two-term split (2)
static uint32_t tc_kd_of(const expann_gpu* h, int mode) {
	const uint32_t seg = (uint32_t)((h->d + 63) / 64 * 64);
	if (mode == 4) return (uint32_t)(2 * h->d_pad); // tf32: the fp32 rows themselves, counted in 2-byte units like the others
	return mode == 2 ? 3 * seg : seg;
}

int launch_filter_pair_tf32(expann_gpu* h, const CUtensorMap& tq, const Tc2Params& p, dim3 grid, cudaStream_t st); // tc_tf32.cu

// format 4 (kind::tf32): no operand copy -- ||x||^2, its three tf32 terms, statistics, tensor maps over the fp32 base rows
static int tc_prepare_tf32(expann_gpu* h) {
	cudaStream_t st = h->stream;
	const size_t n_pad = (h->n + TC_BN - 1) / TC_BN * TC_BN;
	EXPANN_CUDA(h, h->tc_xn.ensure(n_pad * 4));
	EXPANN_CUDA(h, h->tc_xx32.ensure(h->n * 32));
	if (int rc = reset_stats(h, h->tc_stats, st)) return rc;
	tc_norm_tf32_kernel<<<(unsigned)((h->n + 7) / 8), 256, 0, st>>>(h->base, h->d_pad, (uint32_t)h->d, h->n, h->tc_xn.as<float>(),
	                                                               h->tc_xx32.as<float>(), h->cfg.metric == EXPANN_METRIC_IP ? 0 : 1,
	                                                               h->tc_stats.as<TcStats>());
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	TcStats s;
	if (int rc = read_stats(h, h->tc_stats, &s, st)) return rc;
	h->tc_xmin = f32_from_orderable(s.vmin);
	h->tc_xmax = f32_from_orderable(s.vmax);
	h->tc_xnorm_max = std::sqrt(f32_from_orderable(s.nmax));
	h->tc_kd = tc_kd_of(h, 4);
	if (int rc = encode_tensor_map_2d(h, &h->tmap_x32, 0, h->base, h->d_pad, h->n, h->d_pad * 4, 32, TC2_TN)) return rc;
	if (int rc = encode_tensor_map_2d(h, &h->tmap_xx32, 0, h->tc_xx32.p, 8, h->n, 32, 8, TC2_TN, 1)) return rc;
	h->tc_xb.release(); // the bf16 operand of the probed format is not needed any more
	h->tc_xx.release();
	return EXPANN_OK;
}

// materialise the bf16 operand of `fmt` (1 = one bf16 term per value, 2 = two-term split), ||x||^2 and its three-term
// form, the value statistics and the tensor maps
static int tc_prepare(expann_gpu* h, int fmt, TcStats* s_out) {
	cudaStream_t st = h->stream;
	const size_t n_pad = (h->n + TC_BN - 1) / TC_BN * TC_BN;
	const uint32_t kd = tc_kd_of(h, fmt);
	EXPANN_CUDA(h, h->tc_xn.ensure(n_pad * 4));
	EXPANN_CUDA(h, h->tc_xb.ensure(h->n * (size_t)kd * 2));
	if (int rc = reset_stats(h, h->tc_stats, st)) return rc;
	EXPANN_CUDA(h, h->tc_xx.ensure(h->n * 32));
	tc_convert_kernel<<<(unsigned)((h->n + 7) / 8), 256, 0, st>>>(h->base, h->d_pad, (uint32_t)h->d, h->n,
	                                                             h->tc_xb.as<__nv_bfloat16>(), kd, fmt, 0, h->tc_xn.as<float>(),
	                                                             h->tc_stats.as<TcStats>(), 1.f, h->tc_xx.as<__nv_bfloat16>(),
	                                                             h->cfg.metric == EXPANN_METRIC_IP ? 0 : 1);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	TcStats s;
	if (int rc = read_stats(h, h->tc_stats, &s, st)) return rc;
	h->tc_xmin = f32_from_orderable(s.vmin);
	h->tc_xmax = f32_from_orderable(s.vmax);
	h->tc_xnorm_max = std::sqrt(f32_from_orderable(s.nmax));
	h->tc_kd = kd;
	if (s_out) *s_out = s;
	if (n_pad > h->n) {
		tc_fill_kernel<<<(unsigned)((n_pad - h->n + 255) / 256), 256, 0, st>>>(h->tc_xn.as<float>(), h->n, n_pad, nanf(""));
		h->stats.aux_launches++;
	}
	if (h->cfg.metric == EXPANN_METRIC_IP) { // a' = -dot: no norm term
		EXPANN_CUDA(h, cudaMemsetAsync(h->tc_xn.p, 0, h->n * 4, st));
	}
	EXPANN_CUDA(h, cudaStreamSynchronize(st));
	if (int rc = encode_tensor_map_bf16(h, &h->tmap_xb128, h->tc_xb.p, h->tc_kd, h->n, TC2_TN)) return rc;
	if (int rc = encode_tensor_map_bf16(h, &h->tmap_xx, h->tc_xx.p, 16, h->n, TC2_TN, 16)) return rc;
	return encode_tensor_map_bf16(h, &h->tmap_xb, h->tc_xb.p, h->tc_kd, h->n, TC_BN);
}

static int tc_batch2(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st, bool* fell_back);

// build(): decide the operand format from the base values and materialise the bf16 operand + ||x||^2.
//   mode 1  bf16-exact integers whose accumulator stays below 2^24: one GEMM, error bound 0;
//   mode 3  anything else, one bf16 term per value (the SAME operand as mode 1: kd = d, a third of the split's flops)
//           with the slack its rounding error needs; a batch that ends with too many survivors / a full buffer is redone
//           by the SIMT scan and moves the handle to
//   mode 2  the two-term split [qh|qh|ql].[xh|xl|xh] (kd = 3 d, error 2^-14 |q| max|x|) for good (option tc_real = 0
//           starts there).
int tc_finalize(expann_gpu* h) {
	h->tc_mode = 0;
	h->tc_flag_pending = false;
	if (!h->opt_tc || h->cfg.engine != EXPANN_ENGINE_BRUTE_FORCE) return EXPANN_OK;
	const uint32_t seg = (uint32_t)((h->d + 63) / 64 * 64);
	// d too large for a resident query tile: only the streamed-A pair form (tc_filter3.cuh, ASTREAM) can take it
	if (tc_smem_bytes(3 * seg, TC_NSTAGE_SPLIT) > 227 * 1024 && !h->opt_tc_astream) return EXPANN_OK;
	TcStats s;
	if (int rc = tc_prepare(h, 1, &s)) return rc;
	// integer, bf16-exact base whose accumulator stays below 2^24: single GEMM, error bound 0
	// (same bound as tc_check_exact_kernel applies per batch once the queries are known: signed values let the
	// accumulator reach 2 d amax^2 + ||x||^2_max)
	const float amax = std::max(std::fabs(h->tc_xmin), std::fabs(h->tc_xmax));
	const double nmax = (double)h->tc_xnorm_max * h->tc_xnorm_max, dot2 = 2.0 * (double)h->d * amax * amax;
	const double acc_max = h->tc_xmin >= 0.f ? std::max(dot2, nmax) : dot2 + nmax;
	if (!s.inexact && acc_max < 16777216.0) { h->tc_mode = 1; return EXPANN_OK; }
	if (h->opt_tc_real == 1) {
		// Probe: the first rows of the base as a query batch through the one-term filter. Data whose distances are too
		// concentrated for its slack (GIST-like d = 960, u8 values at d = 512: hundreds of survivors per query) is found out
		// here instead of by the first user batch, whose redo by the SIMT scan costs tens of milliseconds.
		h->tc_mode = 3;
		const size_t pm = std::min<size_t>(h->n, 256), pk = std::min<size_t>(h->n, 10);
		DevBuf probe_keys;
		EXPANN_CUDA(h, probe_keys.ensure(pm * pk * 8));
		bool fb = false;
		int rc = tc_batch2(h, h->base, pm, pk, probe_keys.as<uint64_t>(), h->stream, &fb);
		uint32_t redone = 0;
		if (!rc) {
			EXPANN_CUDA(h, cudaStreamSynchronize(h->stream));
			EXPANN_CUDA(h, cudaMemcpy(&redone, h->tc_fbcount.p, 4, cudaMemcpyDeviceToHost));
			EXPANN_CUDA(h, cudaMemset(h->tc_fbcount.p, 0, 4)); // the probe is not a user batch: not counted in tc_fallbacks
		}
		h->tc_flag_pending = false;
		h->stats.tc_launches = 0;
		if (rc) { probe_keys.release(); return rc; }
		if (!redone) { probe_keys.release(); return EXPANN_OK; }
		// next: tf32 on the fp32 rows (slack 5x tighter, twice the MMA time of one bf16 term, no operand copy), probed alike
		if (h->opt_tc_tf32) {
			if (int r2 = tc_prepare_tf32(h)) { probe_keys.release(); return r2; }
			h->tc_mode = 4;
			rc = tc_batch2(h, h->base, pm, pk, probe_keys.as<uint64_t>(), h->stream, &fb);
			redone = 0;
			if (!rc) {
				EXPANN_CUDA(h, cudaStreamSynchronize(h->stream));
				EXPANN_CUDA(h, cudaMemcpy(&redone, h->tc_fbcount.p, 4, cudaMemcpyDeviceToHost));
				EXPANN_CUDA(h, cudaMemset(h->tc_fbcount.p, 0, 4));
			}
			h->tc_flag_pending = false;
			h->stats.tc_launches = 0;
			probe_keys.release();
			if (rc) return rc;
			if (!redone) return EXPANN_OK;
		} else
			probe_keys.release();
	}
	if (int rc = tc_prepare(h, 2, nullptr)) return rc;
	h->tc_mode = 2;
	return EXPANN_OK;
}

// error bound of the operand format, as e = eps_rel * |q| * max|x| + eps_abs on the a' scale (see the header comment)
static void tc_eps(const expann_gpu* h, float* eps_rel, float* eps_abs) {
	*eps_rel = 0.f;
	*eps_abs = 0.f;
	if (h->tc_mode == 1) return;
	const float acc = (float)h->tc_kd / 16.f * (1.0f / 8388608.0f); // fp32 accumulation over kd/16 MMA steps, 2^-23 each
	if (h->tc_mode == 4)
		// tf32: the tensor core reads the fp32 rows' top 19 bits (truncation: 2^-10), the queries were rounded to tf32 (2^-11);
		// the query operand carries the factor -2; K = 8 per step: d_pad / 8 accumulation steps
		*eps_rel = 2.0f * (1.0f / 1024.0f + 1.0f / 2048.0f + 1.0f / 1048576.0f) + (float)h->d_pad / 8.f * (1.0f / 8388608.0f);
	else if (h->tc_mode == 3)
		// one bf16 term per operand: unit roundoff 2^-8 each, so a product is off by at most (2^-7 + 2^-16) |q_i x_i|; the
		// query operand carries the factor -2 (L2; -1 for the inner product, where this is twice what is needed)
		*eps_rel = 2.0f * (1.0f / 128.0f + 1.0f / 65536.0f) + acc;
	else
		// two-term split: 3 * 2^-18 for the dropped terms + fp32 accumulation over kd/16 MMA steps; 2^-14 covers kd <= 1024
		*eps_rel = (1.0f / 16384.0f) * std::max(1.0f, (float)h->tc_kd / 1024.0f);
	*eps_abs = 1e-6f * (h->tc_xnorm_max * h->tc_xnorm_max + 1e-30f); // fp32 rounding of ||x||^2 and of the final fma
}

// one-term format: has an earlier batch been redone? (its flag word was copied to page-locked memory behind the batch;
// looked at here without waiting) -> the handle moves to the two-term split for good
static int tc_adapt_format(expann_gpu* h) {
	if ((h->tc_mode != 3 && h->tc_mode != 4) || !h->tc_flag_pending) return EXPANN_OK;
	if (cudaEventQuery(h->tc_flag_event) != cudaSuccess) { cudaGetLastError(); return EXPANN_OK; }
	h->tc_flag_pending = false;
	if (*h->tc_flag_host.as<uint32_t>() == 0) return EXPANN_OK;
	EXPANN_CUDA(h, cudaDeviceSynchronize()); // operand buffers are about to be rewritten
	if (h->tc_mode == 3 && h->opt_tc_tf32) { // one bf16 term -> tf32 -> two-term split
		if (int rc = tc_prepare_tf32(h)) return rc;
		h->tc_mode = 4;
	} else {
		if (int rc = tc_prepare(h, 2, nullptr)) return rc;
		h->tc_mode = 2;
	}
	h->stats.tc_mode = (uint64_t)h->tc_mode;
	return EXPANN_OK;
}
static int tc_watch_flag(expann_gpu* h, const uint32_t* flags, cudaStream_t st) {
	if ((h->tc_mode != 3 && h->tc_mode != 4) || h->tc_flag_pending) return EXPANN_OK;
	if (!h->tc_flag_event) {
		EXPANN_CUDA(h, h->tc_flag_host.ensure(4));
		EXPANN_CUDA(h, cudaEventCreateWithFlags(&h->tc_flag_event, cudaEventDisableTiming));
	}
	EXPANN_CUDA(h, cudaMemcpyAsync(h->tc_flag_host.p, flags, 4, cudaMemcpyDeviceToHost, st));
	EXPANN_CUDA(h, cudaEventRecord(h->tc_flag_event, st));
	h->tc_flag_pending = true;
	return EXPANN_OK;
}

template <int NSTAGE>
static int launch_filter(expann_gpu* h, const CUtensorMap& tq, const TcParams& p, dim3 grid, cudaStream_t st) {
	const size_t smem = tc_smem_bytes(p.kd, NSTAGE);
	if (int rc = ensure_dyn_smem(h, (const void*)tc_filter_kernel<NSTAGE>, smem)) return rc;
	tc_filter_kernel<NSTAGE><<<grid, TC_THREADS, smem, st>>>(tq, h->tmap_xb, p);
	EXPANN_CUDA(h, cudaGetLastError());
	return EXPANN_OK;
}

static int tc_batch(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st, bool* fell_back) {
	*fell_back = false;
	const uint32_t kd = h->tc_kd;
	const bool ip = h->cfg.metric == EXPANN_METRIC_IP;
	EXPANN_CUDA(h, h->tc_qb.ensure(m * (size_t)kd * 2));
	EXPANN_CUDA(h, h->tc_qn.ensure(m * 4));
	EXPANN_CUDA(h, h->tc_thr.ensure(m * 4));
	EXPANN_CUDA(h, h->tc_cnt.ensure(m * 4 + 16));
	EXPANN_CUDA(h, h->tc_cand.ensure(m * (size_t)TC_CAP * 8));
	uint32_t* flags = h->tc_cnt.as<uint32_t>() + m;
	if (int rc = reset_stats(h, h->tc_qstats, st)) return rc;
	tc_convert_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(d_q, h->d_pad, (uint32_t)h->d, m, h->tc_qb.as<__nv_bfloat16>(), kd,
	                                                          h->tc_mode == 2 ? 2 : 1, 1, h->tc_qn.as<float>(), h->tc_qstats.as<TcStats>());
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	float eps_rel = 0.f, eps_abs = 0.f;
	if (h->tc_mode == 1) {
		// exact only if the queries are bf16-exact integers too and no intermediate can leave the 2^24 integer range
		TcStats s;
		if (int rc = read_stats(h, h->tc_qstats, &s, st)) return rc;
		const float lo = std::min(h->tc_xmin, f32_from_orderable(s.vmin)), hi = std::max(h->tc_xmax, f32_from_orderable(s.vmax));
		const float amax = std::max(std::fabs(lo), std::fabs(hi));
		const double lim = 16777216.0, dd = (double)h->d;
		const bool exact = !s.inexact && amax <= 256.f && dd * (hi - lo) * (hi - lo) < lim && dd * amax * amax < lim;
		if (!exact) { *fell_back = true; return launch_scan_f32_simt(h, d_q, m, k, d_keys_out, st); }
	} else
		tc_eps(h, &eps_rel, &eps_abs);
	CUtensorMap tq;
	if (int rc = encode_tensor_map_bf16(h, &tq, h->tc_qb.p, kd, m, TC_BM)) return rc;
	EXPANN_CUDA(h, cudaMemsetAsync(h->tc_cnt.p, 0, m * 4 + 16, st));
	tc_fill_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(h->tc_thr.as<float>(), 0, m, INFINITY);
	h->stats.aux_launches++;
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
	const uint32_t qblocks = (uint32_t)((m + TC_BM - 1) / TC_BM);
	const uint32_t total_tiles = (uint32_t)((h->n + TC_BN - 1) / TC_BN);
	TcParams p{};
	p.xn = h->tc_xn.as<float>();
	p.thr = h->tc_thr.as<float>();
	p.cand = h->tc_cand.as<uint64_t>();
	p.cnt = h->tc_cnt.as<uint32_t>();
	p.cap = TC_CAP;
	p.m = (uint32_t)m;
	p.n = (uint32_t)h->n;
	p.kd = kd;
	p.scale = ip ? -1.f : -2.f;
	uint32_t tile0 = 0, chunk_tiles = TC_FIRST_ROWS / TC_BN;
	while (tile0 < total_tiles) {
		const uint32_t nt = std::min(chunk_tiles, total_tiles - tile0);
		const bool last = tile0 + nt >= total_tiles;
		// row splits: aim at >= 4 CTAs per SM-slot overall, each CTA with a few tiles to amortise the resident query tile
		uint32_t rsplit = std::max(1u, std::min(nt, (uint32_t)((4u * sms + qblocks - 1) / qblocks)));
		p.tile0 = tile0;
		p.ntiles = nt;
		p.tiles_per_cta = (nt + rsplit - 1) / rsplit;
		rsplit = (nt + p.tiles_per_cta - 1) / p.tiles_per_cta;
		dim3 grid(qblocks, rsplit);
		cudaEvent_t t0 = timing_begin(h, st);
		int rc = h->tc_mode != 2 ? launch_filter<TC_NSTAGE_EXACT>(h, tq, p, grid, st) : launch_filter<TC_NSTAGE_SPLIT>(h, tq, p, grid, st);
		timing_end(h, t0, st);
		if (rc) return rc;
		h->stats.tc_launches++;
		const unsigned sb = (unsigned)((m + 3) / 4);
		if (k <= 32)
			tc_select_kernel<1><<<sb, 128, 0, st>>>(p.cand, p.cnt, TC_CAP, p.m, (uint32_t)k, h->tc_qn.as<float>(), eps_rel, eps_abs,
			                                        h->tc_xnorm_max, h->tc_thr.as<float>(), flags, last ? 1 : 0, TC_PAD);
		else
			tc_select_kernel<4><<<sb, 128, 0, st>>>(p.cand, p.cnt, TC_CAP, p.m, (uint32_t)k, h->tc_qn.as<float>(), eps_rel, eps_abs,
			                                        h->tc_xnorm_max, h->tc_thr.as<float>(), flags, last ? 1 : 0, TC_PAD);
		EXPANN_CUDA(h, cudaGetLastError());
		h->stats.aux_launches++;
		tile0 += nt;
		if (tile0 >= 2 * (TC_FIRST_ROWS / TC_BN)) chunk_tiles *= 2; // 2048, 2048, 4096, 8192, ...
	}
	// K4: exact fp32 keys of the survivors, (key, id) order, truncate to k
	RerankParams r{};
	r.base = h->base;
	r.queries = d_q;
	r.cand = h->tc_cand.as<uint64_t>();
	r.cand_stride = TC_CAP;
	r.out_keys = d_keys_out;
	r.m = (uint32_t)m;
	r.k = (uint32_t)k;
	r.kp = TC_PAD;
	r.d_pad = (uint32_t)h->d_pad;
	r.id_add = (uint32_t)h->cfg.id_offset;
	r.metric = ip ? 1 : 0;
	r.scan_order = 1;
	const size_t smem = ((size_t)h->d_pad * 4 + GS_MAXNB * 8) * 4;
	const unsigned blocks = (unsigned)((m + 3) / 4);
	if (k <= 32) {
		if (int rc = ensure_dyn_smem(h, (const void*)rerank_kernel<1>, smem)) return rc;
		rerank_kernel<1><<<blocks, 128, smem, st>>>(r);
	} else {
		if (int rc = ensure_dyn_smem(h, (const void*)rerank_kernel<4>, smem)) return rc;
		rerank_kernel<4><<<blocks, 128, smem, st>>>(r);
	}
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	// overflow check (needs the flag on the host): redo the batch with the SIMT scan if any buffer overflowed
	// A raised flag (queries not bf16-exact in the exact mode, a candidate buffer overflow, too many survivors) means the
	// batch must be redone by the SIMT scan. No host round trip: the redo is queued behind the flag (ScanParams::gate)
	// and returns at once when the flag is down; the launches that did run are counted in tc_fbcount (stats()).
	if (h->tc_fbcount.cap < 4) {
		EXPANN_CUDA(h, h->tc_fbcount.ensure(4));
		EXPANN_CUDA(h, cudaMemsetAsync(h->tc_fbcount.p, 0, 4, st));
	}
	return launch_scan_f32_simt(h, d_q, m, k, d_keys_out, st, flags, h->tc_fbcount.as<uint32_t>());
}


template <int NQB, int NSTAGE>
static int launch_filter2(expann_gpu* h, const CUtensorMap& tq, const Tc2Params& p, dim3 grid, cudaStream_t st) {
	const size_t smem = Tc2Cfg<NQB>::smem_bytes(p.kd, NSTAGE);
	if (int rc = ensure_dyn_smem(h, (const void*)tc_filter2_kernel<NQB, NSTAGE>, smem)) return rc;
	tc_filter2_kernel<NQB, NSTAGE><<<grid, Tc2Cfg<NQB>::THREADS, smem, st>>>(tq, h->tmap_xb128, h->tmap_xx, p);
	EXPANN_CUDA(h, cudaGetLastError());
	return EXPANN_OK;
}

template <int NSTAGE, bool ASTREAM = false>
static int launch_filter_pair(expann_gpu* h, const CUtensorMap& tq, const Tc2Params& p, dim3 grid, cudaStream_t st) {
	const size_t smem = Tc3Cfg::smem_bytes(p.kd, NSTAGE, ASTREAM);
	if (int rc = ensure_dyn_smem(h, (const void*)tc_filter_pair_kernel<NSTAGE, ASTREAM>, smem)) return rc;
	tc_filter_pair_kernel<NSTAGE, ASTREAM><<<grid, Tc3Cfg::THREADS, smem, st>>>(tq, h->tmap_xb128, h->tmap_xx, p); // __cluster_dims__(2,1,1)
	EXPANN_CUDA(h, cudaGetLastError());
	return EXPANN_OK;
}

// row splits of the second form: whole waves of CTAs (>= 97 % of the SM slots of the last wave used), as few
// splits as that allows (every split warms its own bound up), at least 8 tiles per CTA
uint32_t tc2_pick_rsplit(uint32_t qblocks, uint32_t ntiles, int sms) {
	const uint32_t max_split = std::max(1u, std::min((uint32_t)TC2_MAX_SPLIT, ntiles / 8u));
	uint32_t best = 1;
	double best_eff = 0.0;
	for (uint32_t w = 1; w <= 24; ++w) {
		uint32_t rs = (uint32_t)((uint64_t)sms * w / qblocks);
		if (rs < 1) continue;
		if (rs > max_split) rs = max_split;
		const uint32_t ctas = qblocks * rs;
		const double eff = (double)ctas / (double)((ctas + sms - 1) / sms * sms);
		if (eff > best_eff + 1e-9) { best_eff = eff; best = rs; }
		if (eff >= 0.97 || rs == max_split) break;
	}
	return best;
}

// second form of K3 (k <= 16): ONE filter launch with in-kernel bounds, one select launch, one re-rank launch
static int tc_batch2(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st, bool* fell_back) {
	*fell_back = false;
	if (int rc = tc_adapt_format(h)) return rc;
	const uint32_t kd = h->tc_kd;
	const bool ip = h->cfg.metric == EXPANN_METRIC_IP;
	const int nqb = Tc2Cfg<2>::smem_bytes(kd, TC2_NSTAGE) <= 227 * 1024 && m > TC_BM ? 2 : 1;
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
	// pair form (tc_filter3.cuh): clusters of two CTAs, 256 queries x 256-row tiles per MMA; each CTA contributes two
	// candidate buffers per row split (one per half tile)
	// streamed-A pair form: the only one for operand rows too long for a resident query tile (any batch size then)
	const bool astream = Tc2Cfg<1>::smem_bytes(kd, 4) > 227 * 1024 || h->opt_tc_astream == 2;
	const bool tf32 = h->tc_mode == 4; // pair form only (any batch size)
	const bool pair = tf32 || astream || (h->opt_tc_pair && m > TC_BM && Tc3Cfg::smem_bytes(kd, 4) <= 227 * 1024);
	const uint32_t qblocks = pair ? 2 * (uint32_t)((m + 255) / 256) : (uint32_t)((m + TC_BM * nqb - 1) / (TC_BM * nqb));
	const uint32_t ntiles = (uint32_t)((h->n + (pair ? TC3_TN : TC2_TN) - 1) / (pair ? TC3_TN : TC2_TN));
	// pair form, large bases: a short first launch over the leading boot_tiles tiles seeds every stream of the main launch
	// with the k smallest a' found there (Tc2Params::seed_in): fewer bound-tightening events in the main launch. Worth it
	// only where the epilogue is the limiter (the split mode is MMA-bound: 9.70 ms without, 9.94 with). The seeding launch
	// is ONE wave of clusters: (blocks of 256 queries) x (boot_split row splits of boot_each tiles); its 2 * boot_split
	// lists per query are merged into one by tc_seed_merge_kernel. (Round 1 ran it with one cluster per 256 queries over
	// 64 tiles: ~0.14 ms whatever the batch, break-even at ~8000 queries; split over the idle SMs it pays from small
	// batches on, where a query has the most cold streams -- one GPU's share of C2 at 8 GPUs.)
	const uint32_t qclusters = qblocks / 2u;
	const bool boot_on = pair && (h->opt_tc_boot == 2 || (h->opt_tc_boot == 1 && h->tc_mode != 2)) && ntiles >= 128;
	uint32_t boot_split = boot_on ? std::max(1u, std::min(32u, (uint32_t)(sms / 2) / std::max(1u, qclusters))) : 0u;
	uint32_t boot_each = 0;
	if (boot_on) {
		const uint32_t want_tiles = std::max(8u, std::min((uint32_t)std::max(1, h->opt_tc_boot_tiles), ntiles / 32u));
		boot_each = std::max(2u, (want_tiles + boot_split - 1) / boot_split);
		boot_split = std::max(1u, std::min(boot_split, (want_tiles + boot_each - 1) / boot_each));
	}
	const uint32_t boot_tiles = boot_split * boot_each;
	const uint32_t main_tiles = ntiles - boot_tiles;
	uint32_t rows_split = tc2_pick_rsplit(qblocks, main_tiles, sms);
	if (h->opt_nsplit > 0) rows_split = std::max(1u, std::min({ (uint32_t)h->opt_nsplit, main_tiles, (uint32_t)TC2_MAX_SPLIT / 2u - 1u })); // experiments
	const uint32_t tiles_per_cta = (main_tiles + rows_split - 1) / rows_split;
	rows_split = (main_tiles + tiles_per_cta - 1) / tiles_per_cta;
	const uint32_t rsplit = (pair ? 2 * rows_split : rows_split) + 2u * boot_split; // candidate buffers per query
	EXPANN_CUDA(h, h->tc_qb.ensure(m * (size_t)kd * 2));
	EXPANN_CUDA(h, h->tc_qn.ensure(m * 4));
	EXPANN_CUDA(h, h->tc_gthr.ensure(m * 4));
	if (boot_tiles) EXPANN_CUDA(h, h->tc_seed.ensure(m * (32 + (boot_split > 1 ? 2 * boot_split * 16 : 0)) * 4)); // [m][32] merged | [m][2 * boot_split][16]
	EXPANN_CUDA(h, h->tc_cnt.ensure(m * (size_t)rsplit * 4 + 16));
	EXPANN_CUDA(h, h->tc_cand.ensure(m * (size_t)rsplit * TC2_CAP * 8));
	EXPANN_CUDA(h, h->tc_sel.ensure(m * (size_t)TC_PAD * 8 + 16));
	uint32_t* flags = reinterpret_cast<uint32_t*>(h->tc_sel.as<uint64_t>() + m * (size_t)TC_PAD);
	EXPANN_CUDA(h, h->tc_qstats.ensure(sizeof(TcStats)));
	tc_batch_init_kernel<<<(unsigned)((std::max<size_t>(m, 4) + 255) / 256), 256, 0, st>>>(h->tc_qstats.as<TcStats>(), flags, h->tc_gthr.as<uint32_t>(), (uint32_t)m);
	if (tf32) // the query operand of format 4: fp32 rows, scaled by -2 (-1) and rounded to tf32, [m][d_pad]
		tc_query_tf32_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(d_q, h->d_pad, (uint32_t)h->d, (uint32_t)h->d_pad, m, h->tc_qb.as<float>(),
		                                                             ip ? -1.f : -2.f, h->tc_qn.as<float>());
	else
		tc_convert_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(d_q, h->d_pad, (uint32_t)h->d, m, h->tc_qb.as<__nv_bfloat16>(), kd,
		                                                          h->tc_mode == 2 ? 2 : 1, 1, h->tc_qn.as<float>(), h->tc_qstats.as<TcStats>(), ip ? -1.f : -2.f);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches += 2;
	float eps_rel = 0.f, eps_abs = 0.f;
	if (h->tc_mode == 1) {
		// exact only if the queries are bf16-exact integers too: checked on the device, a violation raises flags[0]
		tc_check_exact_kernel<<<1, 1, 0, st>>>(h->tc_qstats.as<TcStats>(), h->tc_xmin, h->tc_xmax, (float)h->d,
		                                      h->tc_xnorm_max * h->tc_xnorm_max, flags);
		h->stats.aux_launches++;
	} else
		tc_eps(h, &eps_rel, &eps_abs);
	CUtensorMap tq;
	if (tf32) {
		if (int rc = encode_tensor_map_2d(h, &tq, 0, h->tc_qb.p, h->d_pad, m, h->d_pad * 4, 32, TC_BM)) return rc;
	} else if (int rc = encode_tensor_map_bf16(h, &tq, h->tc_qb.p, kd, m, TC_BM))
		return rc;
	Tc2Params p{};
	p.xn = h->tc_xn.as<float>();
	p.qn = h->tc_qn.as<float>();
	p.gthr = h->tc_gthr.as<uint32_t>();
	p.cand = h->tc_cand.as<uint64_t>();
	p.cnt = h->tc_cnt.as<uint32_t>();
	p.cap = TC2_CAP;
	p.m = (uint32_t)m;
	p.n = (uint32_t)h->n;
	p.kd = kd;
	p.k = (uint32_t)k;
	p.ntiles = main_tiles;
	p.tiles_per_cta = tiles_per_cta;
	p.rsplit = rsplit;
	p.tile_first = boot_tiles;
	p.buf_first = 2u * boot_split;
	p.seed_in = boot_tiles ? h->tc_seed.as<float>() : nullptr;
	p.seed_out = nullptr;
	p.scale = ip ? -1.f : -2.f;
	p.eps_rel = eps_rel;
	p.eps_abs = eps_abs;
	p.xnorm_max = h->tc_xnorm_max;
	dim3 grid(qblocks, rows_split);
	cudaEvent_t t0 = timing_begin(h, st);
	// ring depth: 6 slots of 20 KB when they fit beside the resident query tile(s), else 4 (split mode, d up to 128)
	const bool deep = Tc2Cfg<1>::smem_bytes(kd, TC2_NSTAGE) <= 227 * 1024;
	if (boot_tiles) {
		Tc2Params b = p;
		b.ntiles = boot_tiles;
		b.tiles_per_cta = boot_each;
		b.tile_first = 0;
		b.buf_first = 0;
		b.seed_in = nullptr;
		b.seed_out = h->tc_seed.as<float>() + (boot_split > 1 ? m * 32 : 0);
		b.seed_lists = 2u * boot_split;
		const dim3 bgrid(qblocks, boot_split);
		int brc = tf32                                                 ? launch_filter_pair_tf32(h, tq, b, bgrid, st)
		          : astream                                            ? launch_filter_pair<5, true>(h, tq, b, bgrid, st)
		          : Tc3Cfg::smem_bytes(kd, TC2_NSTAGE) <= 227 * 1024 ? launch_filter_pair<TC2_NSTAGE>(h, tq, b, bgrid, st)
		                                                              : launch_filter_pair<4>(h, tq, b, bgrid, st);
		if (brc) return brc;
		h->stats.tc_launches++;
		if (boot_split > 1) {
			tc_seed_merge_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(b.seed_out, b.seed_lists, p.m, p.k, h->tc_seed.as<float>());
			EXPANN_CUDA(h, cudaGetLastError());
			h->stats.aux_launches++;
		}
	}
	int rc = tf32     ? launch_filter_pair_tf32(h, tq, p, grid, st)
	         : astream  ? launch_filter_pair<5, true>(h, tq, p, grid, st)
	         : pair   ? (Tc3Cfg::smem_bytes(kd, TC2_NSTAGE) <= 227 * 1024 ? launch_filter_pair<TC2_NSTAGE>(h, tq, p, grid, st)
	                                                                     : launch_filter_pair<4>(h, tq, p, grid, st))
	         : nqb == 2 ? launch_filter2<2, TC2_NSTAGE>(h, tq, p, grid, st)
	         : deep     ? launch_filter2<1, TC2_NSTAGE>(h, tq, p, grid, st)
	                    : launch_filter2<1, 4>(h, tq, p, grid, st);
	timing_end(h, t0, st);
	if (rc) return rc;
	h->stats.tc_launches++;
	h->stats.last_qtile = pair ? 256 : TC_BM * nqb;
	h->stats.last_nsplit = rows_split;
	if (rsplit > 32) // many cold splits per query: one block per query
		tc_select2_block_kernel<1><<<(unsigned)m, 256, 0, st>>>(p.cand, p.cnt, TC2_CAP, rsplit, p.m, p.k, p.qn, eps_rel, eps_abs,
		                                                        h->tc_xnorm_max, h->tc_sel.as<uint64_t>(), TC_PAD, flags, p.gthr);
	else
		tc_select2_kernel<1><<<(unsigned)((m + 3) / 4), 128, 0, st>>>(p.cand, p.cnt, TC2_CAP, rsplit, p.m, p.k, p.qn, eps_rel, eps_abs,
		                                                             h->tc_xnorm_max, h->tc_sel.as<uint64_t>(), TC_PAD, flags);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	RerankParams r{};
	r.base = h->base;
	r.queries = d_q;
	r.cand = h->tc_sel.as<uint64_t>();
	r.cand_stride = TC_PAD;
	r.out_keys = d_keys_out;
	r.m = (uint32_t)m;
	r.k = (uint32_t)k;
	r.kp = TC_PAD;
	r.d_pad = (uint32_t)h->d_pad;
	r.id_add = (uint32_t)h->cfg.id_offset;
	r.metric = ip ? 1 : 0;
	r.scan_order = 1;
	const size_t smem = ((size_t)h->d_pad * 4 + GS_MAXNB * 8) * 4;
	if (int rc = ensure_dyn_smem(h, (const void*)rerank_kernel<1>, smem)) return rc;
	rerank_kernel<1><<<(unsigned)((m + 3) / 4), 128, smem, st>>>(r);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	// A raised flag (queries not bf16-exact in the exact mode, a candidate buffer overflow, too many survivors) means the
	// batch must be redone by the SIMT scan. No host round trip: the redo is queued behind the flag (ScanParams::gate)
	// and returns at once when the flag is down; the launches that did run are counted in tc_fbcount (stats()).
	if (h->tc_fbcount.cap < 4) {
		EXPANN_CUDA(h, h->tc_fbcount.ensure(4));
		EXPANN_CUDA(h, cudaMemsetAsync(h->tc_fbcount.p, 0, 4, st));
	}
	if (int rc = launch_scan_f32_simt(h, d_q, m, k, d_keys_out, st, flags, h->tc_fbcount.as<uint32_t>())) return rc;
	return tc_watch_flag(h, flags, st);
}

// entry used by the brute-force engine for large batches
bool tc_first_form_fits(const expann_gpu* h) { // (the chunked first form has no tf32 variant)
	return h->tc_mode != 4 && tc_smem_bytes(h->tc_kd, h->tc_mode != 2 ? TC_NSTAGE_EXACT : TC_NSTAGE_SPLIT) <= 227 * 1024;
}

int launch_scan_tc(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st) {
	for (size_t q0 = 0; q0 < m; q0 += TC_QBATCH) {
		const size_t mb = std::min<size_t>(TC_QBATCH, m - q0);
		bool fb = false;
		const bool v2 = k <= TC2_KK && (!h->opt_tc_v1 || !tc_first_form_fits(h));
		if (int rc = v2 ? tc_batch2(h, d_q + q0 * h->d_pad, mb, k, d_keys_out + q0 * k, st, &fb)
		                : tc_batch(h, d_q + q0 * h->d_pad, mb, k, d_keys_out + q0 * k, st, &fb))
			return rc;
		if (!v2 && !fb) { h->stats.last_qtile = TC_BM; h->stats.last_nsplit = 0; }
	}
	return EXPANN_OK;
}

} // namespace expann
