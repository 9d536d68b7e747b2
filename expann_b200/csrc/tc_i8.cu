// The quantised engine's compressed scan on the tensor cores: K6 as a candidate filter with tcgen05.mma.kind::i8.
//
// Reference path: quantizer.h's "score every candidate with the cheap scorer, keep the best" (src/quantizer.h:20-59) with
// dist2_compressed (src/antitopo_engine.h:38-61) on quantizer_simple<uint8_t> codes, or the int8 distance
// (src/distance.h:29-53, sign-corrected) on quantizer_ranged_q8 codes — BASELINE config 5's "quantizer compressed scan +
// fp32 rerank". The SIMT form (quant_scan.cuh, dp4a) is instruction-bound: 22.6 ms per 1000 queries on 1M x 960 codes.
// Here  sum (q-c)^2 - sum q^2 = sum c^2 - 2 sum q c  comes from the pair kernel of K3 (tc_filter3.cuh, KIND = 1): query
// codes x base codes as u8 operands, exact int32 dot products in TMEM (2x the bf16 rate on 4x smaller operands than the
// split-bf16 form), the code norms added in the epilogue; in-kernel bounds, candidate buffers and the select kernels
// are K3's (error bound 0). The k' best codes by (score, id) — bit-identical to the SIMT scan's — go to the fp32 re-rank.
//
// k' > 16 (the engine's default is min(128, 10 k)): a stream keeps only ITS 16 smallest values, so the union of the
// buffers holds the exact top k' only if no stream owned more than 16 of them. tc_select2 verifies exactly that (the
// k'-th smallest key over all buffers must not exceed any stream's final bound, Tc2Params::thr_out); a batch that fails
// (rows ordered so that the best codes pile up in one row range), or overflows a buffer, raises the device flag and is
// redone by the SIMT scan queued behind it (QScanParams::gate) — no host round trip, same result.
#include "engine.h"
#include "tc_filter3.cuh"
#include <algorithm>

namespace expann {


constexpr int32_t CODE_BIAS = 0x40000000; // a'' = bias + sum c^2 - 2 sum q c stays a positive normal float bit pattern
constexpr uint32_t CODE_PAD_NORM = 0x7F000000u;
constexpr uint32_t TCQ_CAP = 512, TCQ_PAD = 256;

static __global__ void code_bias_kernel(const int32_t* __restrict__ cnorm, int32_t* __restrict__ cnb, size_t n, size_t n_pad, int l2) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_pad) return;
	cnb[i] = i < n ? (l2 ? cnorm[i] : 0) + CODE_BIAS : (int32_t)CODE_PAD_NORM;
}

bool tc_code_ready(const expann_gpu* h) { return h->q_tc_ready; }
// k' > 16 needs enough row streams per query (see launch_code_filter): tiny bases stay on the SIMT scan
bool tc_code_fits(const expann_gpu* h, size_t kp) {
	if (kp <= TC2_KK) return true;
	const size_t ntiles = (h->n + TC3_TN - 1) / TC3_TN;
	return 2 * std::min<size_t>(ntiles, (kp + 5) / 6) * TC2_KK >= 3 * kp;
}

// build(): biased code norms and a tensor map with [128 rows][128 code bytes] boxes over the code matrix
int tc_code_prepare(expann_gpu* h) {
	h->q_tc_ready = false;
	if (!h->opt_quant_tc) return EXPANN_OK;
	// 2 * dc * 255^2 must stay below the bias' headroom (2^29)
	if (h->dc_pad > 4096) return EXPANN_OK;
	const size_t n_pad = (h->n + TC3_TN - 1) / TC3_TN * TC3_TN + TC3_TN;
	EXPANN_CUDA(h, h->q_cnb.ensure(n_pad * 4));
	code_bias_kernel<<<(unsigned)((n_pad + 255) / 256), 256, 0, h->stream>>>(h->q_cnorm.as<int32_t>(), h->q_cnb.as<int32_t>(), h->n, n_pad,
	                                                                          h->cfg.metric == EXPANN_METRIC_IP ? 0 : 1);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	EXPANN_CUDA(h, cudaStreamSynchronize(h->stream));
	if (int rc = encode_tensor_map_2d(h, &h->tmap_codes128, 1, h->q_codes.p, h->dc_pad, h->n, h->dc_pad, 128, TC2_TN)) return rc;
	h->q_tc_ready = true;
	return EXPANN_OK;
}

template <int NSTAGE, bool ASTREAM>
static int launch_code_pair(expann_gpu* h, const CUtensorMap& tq, const Tc2Params& p, dim3 grid, cudaStream_t st) {
	const size_t smem = Tc3Cfg::smem_bytes(p.kd, NSTAGE, ASTREAM);
	if (int rc = ensure_dyn_smem(h, (const void*)tc_filter_pair_kernel<NSTAGE, ASTREAM, 1>, smem)) return rc;
	// the norm-term tensor map of the float path is not used by KIND 1: the code map stands in
	tc_filter_pair_kernel<NSTAGE, ASTREAM, 1><<<grid, Tc3Cfg::THREADS, smem, st>>>(tq, h->tmap_codes128, h->tmap_codes128, p); // __cluster_dims__(2,1,1)
	EXPANN_CUDA(h, cudaGetLastError());
	return EXPANN_OK;
}

// query codes (h->qq_codes, [m][dc_pad]) x base codes -> the kp best codes per query by (score, id) in d_cand_out [m][kp]
// (keys ordered like the SIMT scan's; low 32 bits = local row). *flags_out: device word that is non-zero when the batch
// must be redone.
int launch_code_filter(expann_gpu* h, size_t m, size_t kp, uint64_t* d_cand_out, cudaStream_t st, const uint32_t** flags_out) {
	const bool ip = h->cfg.metric == EXPANN_METRIC_IP;
	const uint32_t kd = (uint32_t)(h->dc_pad / 2); // 2-byte units, so that a k-block is 64 of them = 128 code bytes
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
	const bool astream = Tc3Cfg::smem_bytes(kd, 4) > 227 * 1024;
	const uint32_t qblocks = 2 * (uint32_t)((m + 255) / 256);
	const uint32_t ntiles = (uint32_t)((h->n + TC3_TN - 1) / TC3_TN);
	uint32_t rows_split = tc2_pick_rsplit(qblocks, ntiles, sms);
	// k' > 16: the union of the streams' 16 smallest must cover the k' smallest overall: at least k' / 3 streams per query
	// (two per row split), i.e. a stream's 16th smallest sits ~5x above the k'-th smallest of all rows
	if (kp > TC2_KK) rows_split = std::max(rows_split, std::min(ntiles, (uint32_t)((kp + 5) / 6)));
	if (h->opt_nsplit > 0) rows_split = std::max(1u, std::min({ (uint32_t)h->opt_nsplit, ntiles, (uint32_t)TC2_MAX_SPLIT / 2u - 1u })); // experiments
	const uint32_t tiles_per_cta = (ntiles + rows_split - 1) / rows_split;
	rows_split = (ntiles + tiles_per_cta - 1) / tiles_per_cta;
	const uint32_t rsplit = 2 * rows_split; // candidate buffers per query: one per (row split, half tile)
	const uint32_t kf = (uint32_t)std::min<size_t>(kp, TC2_KK);
	EXPANN_CUDA(h, h->tc_gthr.ensure(m * 4));
	EXPANN_CUDA(h, h->tc_cnt.ensure(m * (size_t)rsplit * 4 + 16));
	EXPANN_CUDA(h, h->tc_cand.ensure(m * (size_t)rsplit * TCQ_CAP * 8));
	EXPANN_CUDA(h, h->tc_sel.ensure(m * (size_t)TCQ_PAD * 8 + 16));
	EXPANN_CUDA(h, h->tc_qstats.ensure(sizeof(TcStats)));
	if (kp > TC2_KK) EXPANN_CUDA(h, h->q_thrfin.ensure(m * (size_t)rsplit * 4));
	if (h->q_zero.cap < m * 4) { // ||q|| of the float path's error bound: not used here (error 0), but read by the select kernels
		EXPANN_CUDA(h, h->q_zero.ensure(m * 4));
		EXPANN_CUDA(h, cudaMemsetAsync(h->q_zero.p, 0, h->q_zero.cap, st));
	}
	uint32_t* flags = reinterpret_cast<uint32_t*>(h->tc_sel.as<uint64_t>() + m * (size_t)TCQ_PAD);
	tc_batch_init_kernel<<<(unsigned)((std::max<size_t>(m, 4) + 255) / 256), 256, 0, st>>>(h->tc_qstats.as<TcStats>(), flags, h->tc_gthr.as<uint32_t>(), (uint32_t)m);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	CUtensorMap tq;
	if (int rc = encode_tensor_map_2d(h, &tq, 1, h->qq_codes.p, h->dc_pad, m, h->dc_pad, 128, TC_BM)) return rc;
	Tc2Params p{};
	p.qn = h->q_zero.as<float>();
	p.gthr = h->tc_gthr.as<uint32_t>();
	p.cand = h->tc_cand.as<uint64_t>();
	p.cnt = h->tc_cnt.as<uint32_t>();
	p.cap = TCQ_CAP;
	p.m = (uint32_t)m;
	p.n = (uint32_t)h->n;
	p.kd = kd;
	p.k = kf;
	p.ntiles = ntiles;
	p.tiles_per_cta = tiles_per_cta;
	p.rsplit = rsplit;
	p.scale = ip ? -1.f : -2.f;
	p.thr_out = kp > TC2_KK ? h->q_thrfin.as<uint32_t>() : nullptr;
	p.cnb = h->q_cnb.as<int32_t>();
	p.ibias = CODE_BIAS;
	dim3 grid(qblocks, rows_split);
	cudaEvent_t t0 = timing_begin(h, st);
	int rc = astream                                               ? launch_code_pair<5, true>(h, tq, p, grid, st)
	         : Tc3Cfg::smem_bytes(kd, 6) <= 227 * 1024 ? launch_code_pair<6, false>(h, tq, p, grid, st)
	                                                                 : launch_code_pair<4, false>(h, tq, p, grid, st);
	timing_end(h, t0, st);
	if (rc) return rc;
	h->stats.tc_launches++;
	h->stats.last_qtile = 256;
	h->stats.last_nsplit = rows_split;
	const float* qn = h->q_zero.as<float>();
	const uint32_t* thr_fin = p.thr_out;
	if (rsplit > 32) { // many cold splits per query: one block per query (the shared-bound fast path only for k' <= 16)
		const uint32_t* gthr = kp > TC2_KK ? nullptr : p.gthr;
		if (kp <= 32)
			tc_select2_block_kernel<1><<<(unsigned)m, 256, 0, st>>>(p.cand, p.cnt, TCQ_CAP, rsplit, p.m, (uint32_t)kp, qn, 0.f, 0.f, 0.f, h->tc_sel.as<uint64_t>(), TCQ_PAD, flags, gthr, thr_fin);
		else
			tc_select2_block_kernel<4><<<(unsigned)m, 256, 0, st>>>(p.cand, p.cnt, TCQ_CAP, rsplit, p.m, (uint32_t)kp, qn, 0.f, 0.f, 0.f, h->tc_sel.as<uint64_t>(), TCQ_PAD, flags, gthr, thr_fin);
	} else if (kp <= 32)
		tc_select2_kernel<1><<<(unsigned)((m + 3) / 4), 128, 0, st>>>(p.cand, p.cnt, TCQ_CAP, rsplit, p.m, (uint32_t)kp, qn, 0.f, 0.f, 0.f, h->tc_sel.as<uint64_t>(), TCQ_PAD, flags, thr_fin);
	else
		tc_select2_kernel<4><<<(unsigned)((m + 3) / 4), 128, 0, st>>>(p.cand, p.cnt, TCQ_CAP, rsplit, p.m, (uint32_t)kp, qn, 0.f, 0.f, 0.f, h->tc_sel.as<uint64_t>(), TCQ_PAD, flags, thr_fin);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	// the survivors are every key at or below the k'-th smallest (ties included): K7 orders them by (score, id) and keeps k'
	if (int r2 = launch_merge(h, h->tc_sel.as<uint64_t>(), m, 1, TCQ_PAD, TCQ_PAD, 0, kp, 0, d_cand_out, st)) return r2;
	*flags_out = flags;
	return EXPANN_OK;
}

} // namespace expann
