// Host side of K6 (compressed scan) + K4 (fp32 rerank) + quantizer builds + pair-distance primitives.
#include "engine.h"
#include "quant_scan.cuh"
#include "rerank.cuh"
#include <algorithm>

namespace expann {

constexpr int QS_TR = 2, QS_NWARP = 8, QS_STAGES = 3;
constexpr int QS_ROWS = 32 * QS_TR * QS_NWARP;


static int quantize_rows(expann_gpu* h, const float* src, size_t src_stride, size_t n, uint8_t* dst, int32_t* norm,
                         cudaStream_t st) {
	if (n == 0) return EXPANN_OK;
	const unsigned blocks = (unsigned)((n + 7) / 8);
	if (h->quantizer == 0)
		quant_simple_u8_kernel<<<blocks, 256, 0, st>>>(src, src_stride, (uint32_t)h->d, dst, (uint32_t)h->dc_pad, norm, n);
	else
		quant_ranged_q8_kernel<<<blocks, 256, 0, st>>>(src, src_stride, (uint32_t)h->d, dst, (uint32_t)h->dc_pad, norm, n,
		                                              h->q_scale, h->q_offset);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	return EXPANN_OK;
}

// global min / max over all components of the stored rows (quantizer_ranged_q8::build, src/quantizer.h:206-216)
int quant_minmax(expann_gpu* h, float* mn_out, float* mx_out) {
	cudaStream_t st = h->stream;
	EXPANN_CUDA(h, h->q_mm.ensure(8));
	const uint32_t init[2] = { f32_orderable(1.17549435e-38f), f32_orderable(3.402823466e+38f) };
	EXPANN_CUDA(h, cudaMemcpyAsync(h->q_mm.p, init, 8, cudaMemcpyHostToDevice, st));
	minmax_kernel<<<1184, 256, 0, st>>>(h->base, h->d_pad, (uint32_t)h->d, h->n, h->q_mm.as<float>());
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	uint32_t mm[2];
	EXPANN_CUDA(h, cudaMemcpyAsync(mm, h->q_mm.p, 8, cudaMemcpyDeviceToHost, st));
	EXPANN_CUDA(h, cudaStreamSynchronize(st));
	*mx_out = f32_from_orderable(mm[0]);
	*mn_out = f32_from_orderable(mm[1]);
	return EXPANN_OK;
}
// scale_factor / offset from the value range (src/quantizer.h:217-227). A multi-device handle fits ONE range over all
// shards and hands it to every shard (q_range_fixed), so the codes do not depend on the partitioning.
void quant_set_range(expann_gpu* h, float mn, float mx) {
	h->q_scale = 128.0f / (mx - mn);          // :219, q_range() = 128
	h->q_offset = -h->q_scale * mn - 0.0f;    // :220, q_min = 0
	h->q_range_fixed = true;
}

int quant_finalize(expann_gpu* h) {
	cudaStream_t st = h->stream;
	h->dc_pad = (h->d + 127) / 128 * 128;
	EXPANN_CUDA(h, h->q_codes.ensure(h->n * h->dc_pad));
	EXPANN_CUDA(h, h->q_cnorm.ensure(h->n * 4));
	if (h->quantizer == 1 && !h->q_range_fixed) {
		// quantizer_ranged_q8::build (src/quantizer.h:206-227)
		float mn, mx;
		if (int rc = quant_minmax(h, &mn, &mx)) return rc;
		quant_set_range(h, mn, mx);
		h->q_range_fixed = false; // fitted from this handle's own rows: refit on the next build()
	}
	if (int rc = quantize_rows(h, h->base, h->d_pad, h->n, h->q_codes.as<uint8_t>(), h->q_cnorm.as<int32_t>(), st)) return rc;
	EXPANN_CUDA(h, cudaStreamSynchronize(st));
	if (int rc = encode_tensor_map_2d(h, &h->tmap_codes, 1, h->q_codes.p, h->dc_pad, h->n, h->dc_pad, 128, 256)) return rc;
	return tc_code_prepare(h); // operands of the tensor-core form of the scan (tc_i8.cu)
}

template <int QT, int METRIC, int LPL>
static int launch_u8_inst(expann_gpu* h, const QScanParams& p, dim3 grid, cudaStream_t st) {
	using Cfg = ScanCfg<QT, QS_TR, QS_NWARP, QS_STAGES, METRIC, LPL>;
	auto kern = scan_topk_u8_kernel<QT, QS_TR, QS_NWARP, QS_STAGES, METRIC, LPL>;
	size_t merge = (size_t)QS_NWARP * QT * Cfg::KMAX * 8, stg = (size_t)QS_STAGES * Cfg::STAGE_BYTES;
	const size_t smem = 1024 + std::max(merge, stg) + (((size_t)QT * p.dc_pad + 15) & ~(size_t)15) + 2 * QS_STAGES * 8 + 64;
	if (smem > 227 * 1024) return fail(h, EXPANN_ERR_UNSUPPORTED, "quant scan: d too large");
	if (int rc = ensure_dyn_smem(h, (const void*)kern, smem)) return rc;
	kern<<<grid, Cfg::THREADS, smem, st>>>(h->tmap_codes, p);
	EXPANN_CUDA(h, cudaGetLastError());
	return EXPANN_OK;
}
template <int QT, int METRIC> static int launch_u8_k(expann_gpu* h, const QScanParams& p, dim3 grid, cudaStream_t st) {
	if (p.k <= 32) return launch_u8_inst<QT, METRIC, 1>(h, p, grid, st);
	return launch_u8_inst<QT, METRIC, 4>(h, p, grid, st);
}

// queries (fp32, padded) -> codes + norms -> K6 -> K7 -> K4
int launch_quant_scan(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st) {
	size_t kp = h->opt_kprime > 0 ? (size_t)h->opt_kprime : std::min<size_t>(128, std::max<size_t>(k, 10 * k));
	kp = std::max(kp, k);
	if (kp > 128) return fail(h, EXPANN_ERR_UNSUPPORTED, "quant scan: k' above 128");
	EXPANN_CUDA(h, h->qq_codes.ensure(m * h->dc_pad));
	EXPANN_CUDA(h, h->qq_norm.ensure(m * 4));
	EXPANN_CUDA(h, h->q_cand.ensure(m * kp * 8));
	if (int rc = quantize_rows(h, d_q, h->d_pad, m, h->qq_codes.as<uint8_t>(), h->qq_norm.as<int32_t>(), st)) return rc;

	// Large batches: the code scan runs as a tensor-core candidate filter (tc_i8.cu, tcgen05.mma.kind::i8) that leaves the
	// same k' best codes in q_cand; the SIMT scan below is then only its redo, queued behind the filter's device flag.
	const uint32_t* gate = nullptr;
	uint32_t* gate_count = nullptr;
	if (tc_code_ready(h) && h->opt_quant_tc && m >= (size_t)h->opt_quant_tc_min_batch && tc_code_fits(h, kp)) {
		if (int rc = launch_code_filter(h, m, kp, h->q_cand.as<uint64_t>(), st, &gate)) return rc;
		if (h->tc_fbcount.cap < 4) {
			EXPANN_CUDA(h, h->tc_fbcount.ensure(4));
			EXPANN_CUDA(h, cudaMemsetAsync(h->tc_fbcount.p, 0, 4, st));
		}
		gate_count = h->tc_fbcount.as<uint32_t>();
	}

	int qt = h->cfg.qtile;
	if (qt == 0) qt = m >= 8 ? 8 : (m >= 3 ? 4 : 1);
	const uint32_t qtiles = (uint32_t)((m + qt - 1) / qt);
	const uint32_t groups = (uint32_t)((h->n + QS_ROWS - 1) / QS_ROWS);
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
	// whole waves of CTAs (1 CTA per SM), as few row splits as fill them: with k' ~ 100 every split pays a long list
	// warm-up (the per-warp lists stay unselective for thousands of rows), so 125 query tiles x 1 split (22.6 ms at C5)
	// beat 125 x 19 (41.4 ms); a single query tile still gets one split per SM
	const uint32_t waves = std::max(1u, (qtiles + (uint32_t)sms - 1) / (uint32_t)sms);
	const uint32_t want = std::max(1u, (uint32_t)((uint64_t)sms * waves / qtiles));
	uint32_t nsplit = std::max(1u, std::min(want, std::max(1u, groups / 8u)));
	if (h->opt_nsplit > 0) nsplit = std::min((uint32_t)h->opt_nsplit, groups);
	const uint32_t gps = (groups + nsplit - 1) / nsplit;
	nsplit = (groups + gps - 1) / gps;
	QScanParams p{};
	p.qcodes = h->qq_codes.as<uint8_t>();
	p.qnorm = h->qq_norm.as<int32_t>();
	p.cnorm = h->q_cnorm.as<int32_t>();
	p.m = (uint32_t)m;
	p.k = (uint32_t)kp;
	p.dc_pad = (uint32_t)h->dc_pad;
	p.n = (uint32_t)h->n;
	p.rows_per_split = gps * QS_ROWS;
	p.nsplit = nsplit;
	p.gate = gate;
	p.gate_count = gate_count;
	uint64_t* scan_out = h->q_cand.as<uint64_t>();
	if (nsplit > 1) {
		EXPANN_CUDA(h, h->partial.ensure(m * nsplit * kp * 8));
		scan_out = h->partial.as<uint64_t>();
	}
	p.partial = scan_out;
	dim3 grid(qtiles, nsplit);
	const bool ip = h->cfg.metric == EXPANN_METRIC_IP;
	cudaEvent_t t0 = gate ? nullptr : timing_begin(h, st);
	int rc;
	if (qt == 8) rc = ip ? launch_u8_k<8, METRIC_IP>(h, p, grid, st) : launch_u8_k<8, METRIC_L2>(h, p, grid, st);
	else if (qt == 4) rc = ip ? launch_u8_k<4, METRIC_IP>(h, p, grid, st) : launch_u8_k<4, METRIC_L2>(h, p, grid, st);
	else rc = ip ? launch_u8_k<1, METRIC_IP>(h, p, grid, st) : launch_u8_k<1, METRIC_L2>(h, p, grid, st);
	timing_end(h, t0, st);
	if (rc) return rc;
	if (gate)
		h->stats.aux_launches++; // the queued redo: returns at once while the flag is down
	else {
		h->stats.scan_launches++;
		h->stats.last_nsplit = nsplit;
		h->stats.last_qtile = qt;
	}
	if (nsplit > 1)
		if (int r2 = launch_merge(h, scan_out, m, nsplit, kp, (size_t)nsplit * kp, kp, kp, 0, h->q_cand.as<uint64_t>(), st, gate)) return r2;
	// K4: exact fp32 keys of the k' survivors, final (key, id) order, truncation to k
	RerankParams r{};
	r.base = h->base;
	r.queries = d_q;
	r.cand = h->q_cand.as<uint64_t>();
	r.out_keys = d_keys_out;
	r.m = (uint32_t)m;
	r.k = (uint32_t)k;
	r.kp = (uint32_t)kp;
	r.cand_stride = (uint32_t)kp;
	r.d_pad = (uint32_t)h->d_pad;
	r.id_add = (uint32_t)h->cfg.id_offset;
	r.metric = ip ? 1 : 0;
	const size_t smem = ((size_t)h->d_pad * 4 + GS_MAXNB * 8) * 4;
	const unsigned blocks = (unsigned)((m + 3) / 4);
	if (kp <= 32) {
		if (int rc = ensure_dyn_smem(h, (const void*)rerank_kernel<1>, smem)) return rc;
		rerank_kernel<1><<<blocks, 128, smem, st>>>(r);
	} else {
		if (int rc = ensure_dyn_smem(h, (const void*)rerank_kernel<4>, smem)) return rc;
		rerank_kernel<4><<<blocks, 128, smem, st>>>(r);
	}
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	return EXPANN_OK;
}

} // namespace expann

using namespace expann;

extern "C" int expann_gpu_pair_distances(int device, int kind, const void* a, const void* b, size_t npairs, size_t d,
                                         float* out_f, int32_t* out_i) {
	if (!a || !b || kind < 0 || kind > 4 || (kind <= 1 ? !out_f : !out_i)) return EXPANN_ERR_ARG;
	if (npairs == 0) return EXPANN_OK;
	int prev = -1;
	cudaGetDevice(&prev);
	if (device >= 0 && cudaSetDevice(device) != cudaSuccess) return EXPANN_ERR_CUDA;
	const size_t ea = kind <= 2 ? 4 : (kind == 3 ? 1 : 2), eb = kind <= 1 ? 4 : (kind == 4 ? 2 : 1);
	void *da = nullptr, *db = nullptr, *dout = nullptr;
	int rc = EXPANN_OK;
	if (cudaMalloc(&da, npairs * d * ea) != cudaSuccess || cudaMalloc(&db, npairs * d * eb) != cudaSuccess ||
	    cudaMalloc(&dout, npairs * 4) != cudaSuccess)
		rc = EXPANN_ERR_CUDA;
	if (!rc && (cudaMemcpy(da, a, npairs * d * ea, cudaMemcpyHostToDevice) != cudaSuccess ||
	            cudaMemcpy(db, b, npairs * d * eb, cudaMemcpyHostToDevice) != cudaSuccess))
		rc = EXPANN_ERR_CUDA;
	if (!rc) {
		pair_distance_kernel<<<(unsigned)((npairs + 3) / 4), 128>>>(kind, da, db, npairs, (uint32_t)d, (float*)dout, (int32_t*)dout);
		if (cudaMemcpy(kind <= 1 ? (void*)out_f : (void*)out_i, dout, npairs * 4, cudaMemcpyDeviceToHost) != cudaSuccess)
			rc = EXPANN_ERR_CUDA;
	}
	cudaFree(da);
	cudaFree(db);
	cudaFree(dout);
	if (prev >= 0) cudaSetDevice(prev);
	return rc;
}
