// Decode engine: the whole per-token step of the reference's DirectBitNetInference.forward
// (scripts/direct_inference_full.py:267-302) as one CUDA-graph-captured sequence of our kernels, with
// resident tile-layout weights, a paged KV pool and a free-list page allocator
// (KVCacheManager semantics, tests/test_kv_cache_cpp.py:44-70).  Batched: every sequence is an
// independent copy of the reference's bs=1 loop (per-token activation scales).
//
// Per layer (11 launches):  rmsnorm+quant -> qkv linear -> rope+kv append -> paged attention (+merge)
//   -> sub-norm+quant -> o linear (+residual) -> rmsnorm+quant -> gate|up linear (relu^2*up)
//   -> sub-norm+quant -> down linear (+residual).
#include "kernels.cuh"
#include "fused_common.cuh"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

using namespace wf;

namespace {

struct LayerW {
    uint8_t *qkv = nullptr, *o = nullptr, *gu = nullptr, *down = nullptr;   // b200_tile_v1
    float* scales = nullptr;      // device [7]: q,k,v,o,gate,up,down
    float h_scales[7] = {1, 1, 1, 1, 1, 1, 1};
    __nv_bfloat16 *g_in = nullptr, *g_post = nullptr, *g_attn = nullptr, *g_ffn = nullptr;
    uint8_t* stage[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // I2_S staging until finalize
};

const char* kLinNames[7] = {"q", "k", "v", "o", "gate", "up", "down"};

}  // namespace

struct wf_engine {
    wf_model_config cfg;
    cudaStream_t stream = nullptr;
    std::vector<LayerW> layers;
    __nv_bfloat16 *embed = nullptr, *final_norm = nullptr, *cos_tab = nullptr, *sin_tab = nullptr;
    // local (per TP rank) dims: H / Hkv heads, ffn rows, attn_dim = H*D of THIS rank; *_full = the whole model's
    int H, Hkv, D, hidden, ffn, qkv_rows, attn_dim;
    int attn_full, ffn_full;
    // tensor parallel state (tp.size == 1: everything below degenerates to local buffers, no flags)
    TpComm tp;
    unsigned char* comm = nullptr;           // this rank's exchange buffer (see tp_comm.cuh)
    size_t comm_bytes = 0, off_attn = 0, off_h = 0, off_acc_o = 0, off_acc_down = 0, off_cand = 0, off_flags = 0;
    void* peer_opened[TP_MAX] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool tp_attached = false;
    int* d_step = nullptr;
    int* d_err = nullptr;                    // device error word (bounded spins that gave up), mirrored into h_err_pinned at every host sync
    int* h_err_pinned = nullptr;
    __nv_bfloat16* h_full = nullptr;
    int v_lo = 0, v_n = 0;                   // this rank's vocab slice of the tied LM head
    unsigned long long* trace = nullptr;     // WF_TRACE=1: 16 globaltimer stamps per linear_small launch
    // kv
    __nv_bfloat16* pool = nullptr;   // bf16 pool, or the int8 region of an int8 pool (kv_scales != null)
    float* kv_scales = nullptr;      // int8 KV: scale region [page][layer][K|V][head][slot]
    int max_blocks = 0;
    int32_t* d_block_table = nullptr;
    std::vector<int32_t> h_block_table;
    std::vector<int32_t> free_pages;
    std::vector<int32_t> h_pos;
    bool table_dirty = true;
    int32_t *d_pos = nullptr, *d_len = nullptr, *d_tok = nullptr, *d_next = nullptr;
    int32_t *h_tok_pinned = nullptr, *h_next_pinned = nullptr, *h_pos_pinned = nullptr;
    // activations
    __nv_bfloat16 *hid = nullptr, *qkv = nullptr, *qrot = nullptr, *attn = nullptr, *gu = nullptr, *normed = nullptr;
    int8_t* q8 = nullptr;
    int8_t* q8c = nullptr;   // the same rows in the batched tail kernel's chunk-major image (QMAP_CHUNKED, fused_common.cuh)
    float* a_scale = nullptr;
    int32_t* rowsum = nullptr;
    float* logits = nullptr;
    void *ws_lin = nullptr, *ws_attn = nullptr, *ws_lm = nullptr;
    size_t ws_lin_bytes = 0, ws_attn_bytes = 0, ws_lm_bytes = 0;
    size_t dev_bytes = 0;
    bool finalized = false;
    bool pdl = false;
    int linear_algo = 0;
    int hx = 1;              // relu^2(gate)*up in its own small kernel instead of every down-proj CTA's prologue (WF_HX=0 to fold it back)
    int decode_tc = 1;       // batched decode linears (B >= decode_tc_min, single GPU) on tcgen05 split-K REDs (gemm_tc.cu launch_gemm_tc_red) instead of the mma.sync one-shot
                             // GEMM: bit-identical; with the round-2 issue loop / epilogue 3.65 -> 3.23 ms at batch 40 / 2K, 5.48 -> 5.32 ms at batch 64 / 4K (WF_DECODE_TC=0: off)
    int decode_tc_min = 33;  // WF_DECODE_TC_MIN
    int decode_stream = 1;   // WF_DECODE_STREAM: streaming batched GEMM (gemm_stream.cu); 0 off, 1 = batches up to decode_stream_max rows, 2 = every batch it takes (<= 32)
    int decode_stream_mask = 15;  // WF_DECODE_STREAM_MASK: which linears may take it (1 q|k|v, 2 o, 4 gate|up, 8 down) — for bisecting the TP mismatch
    int decode_stream_max = 16;   // WF_DECODE_STREAM_MAX.  Measured (profiles/r02ap_summary.txt), ms per step stream / default: 2B bs2 1.33 / 1.58, bs8 1.43 / 1.63,
                                  // bs16/2K 1.99 / 2.17; 8B bs4 1.97 / 3.00, bs8 2.26 / 3.24, bs16/2K 3.06 / 3.85; 70B bs16/8K 20.0 / 22.9; but 17..32 rows (four n-tiles,
                                  // one CTA per SM): 2B bs32/2K 2.74 / 2.56, 8B bs32/2K 4.55 / 4.20
    int tail_batch = 0;      // WF_TAIL_BATCH=1: the same for 8 < B <= 64 (layer_tail_batch.cu).  Opt-in: measured 5.74 ms vs 5.44 ms per bs64/4K step for the phase path (profiles/r02d_*)
    int small_max = 1;       // WF_SMALL_MAX: largest batch on the GEMV-style kernels (linear_small / layer_tail: the row glue runs per CTA and per row,
                             // ~3 us x B x 4 phases per layer); above it the batched kernels (row glue once per row in its own CTA + batched GEMM).
                             // Measured (profiles/r02j_summary.txt, 2B, 512 ctx): B = 2 / 4 / 8: 1.81 / 2.46 / 3.70 ms on the GEMV-style kernels, 1.72 / 1.74 / 1.75 ms batched
    int tail_lam = 1;        // WF_TAIL_LAM: batch-1 tail kernel hands its phases over flag-in-data (no grid barriers); 0 = grid barriers
    int tail = 1;            // 1: persistent layer-tail kernel (layer_tail.cu) for B <= 8 on single-GPU engines: 2 launches per layer (WF_TAIL=0: phase kernels)
    // grid-barrier arrival counters, one set per launch site, and the epoch (steps so far) of each kernel family:
    // [0] layer_tail_kernel (B <= 8, LT_MAX_PHASES slots per site), [1] layer_tail_batch_kernel (B > 8, 2 * LT_MAX_PHASES slots)
    unsigned long long* tail_bar[2] = {nullptr, nullptr};
    int* tail_epoch = nullptr;   // [2]
    int fused = 1;           // 1: phase kernels (linear_small / glue_rows + gemm_oneshot / fused attention); 0: one kernel per op
    // fused path state: raw fp32 accumulators of the four linears, their per-token activation scale / Sigma a,
    // a second hidden buffer (residual ping-pong) and the attention merge counters
    float *acc_qkv = nullptr, *acc_o = nullptr, *acc_gu = nullptr, *acc_down = nullptr;
    float* lin_asc = nullptr;      // [4][max_batch]
    int32_t* lin_rs = nullptr;     // [4][max_batch]
    __nv_bfloat16* hid2 = nullptr;
    int* merge_cnt = nullptr;
    int launches = 0;
    int last_family = -1;    // step form of the previous step: 0 per-op kernels, 1 tail kernels, 2 phase kernels (see step_graph)
    int attn_seq_hint = 0;   // max sequence length the chunking is sized for (fixed per captured graph)
    cudaGraphExec_t graph_exec[2] = {nullptr, nullptr};   // [feed_back]
    int graph_B[2] = {0, 0};
    int graph_hint[2] = {0, 0};
    std::vector<void*> allocs;
};

namespace {

template <typename T>
int dalloc(wf_engine* e, T** p, size_t count) {
    void* q = nullptr;
    size_t bytes = count * sizeof(T);
    if (bytes == 0) bytes = 16;
    if (cuda_check(cudaMalloc(&q, bytes), "cudaMalloc")) return WF_ERR_CUDA;
    e->allocs.push_back(q);
    e->dev_bytes += bytes;
    *p = (T*)q;
    return WF_OK;
}

int lin_index(const char* name) {
    for (int i = 0; i < 7; ++i)
        if (!strcmp(name, kLinNames[i])) return i;
    return -1;
}

void lin_shape(const wf_engine* e, int idx, int* N, int* K) {
    switch (idx) {
        case 0: *N = e->H * e->D; *K = e->hidden; break;
        case 1: case 2: *N = e->Hkv * e->D; *K = e->hidden; break;
        case 3: *N = e->hidden; *K = e->attn_dim; break;
        case 4: case 5: *N = e->ffn; *K = e->hidden; break;
        default: *N = e->hidden; *K = e->ffn; break;
    }
}

void invalidate_graphs(wf_engine* e) {
    for (int i = 0; i < 2; ++i)
        if (e->graph_exec[i]) { cudaGraphExecDestroy(e->graph_exec[i]); e->graph_exec[i] = nullptr; }
}

int linear(wf_engine* e, const uint8_t* w, int N, int K, int B, const float* w_scale, int nseg, const int* seg_end, int flags,
           void* out, int ldo, const void* residual) {
    wf_linear_args a;
    memset(&a, 0, sizeof a);
    a.w = w; a.w_layout = WF_W_TILE; a.N = N; a.K = K;
    a.q = e->q8; a.M = B; a.ldq = K;
    a.rowsum = e->rowsum; a.act_scale = e->a_scale; a.act_scale_per_token = 1;
    a.w_scale = w_scale; a.nseg = nseg;
    for (int i = 0; i < nseg && i < 4; ++i) a.seg_end[i] = seg_end[i];
    a.out_mode = WF_OUT_LINEAR; a.epi_flags = flags; a.out_dtype = WF_BF16;
    a.out = out; a.ldo = ldo; a.residual = residual;
    a.workspace = e->ws_lin; a.workspace_bytes = e->ws_lin_bytes;
    EpiParams ep;
    int rc = fill_epi(&a, &ep);
    if (rc) return rc;
    int algo = e->linear_algo;
    if (algo == 0) algo = (B <= 8 && gemv_stream_fits(B, K)) ? 1 : 2;
    if (algo == 1 && B <= 8) return launch_gemv_stream(&a, ep, e->stream, e->pdl);
    if (algo == 3) return launch_gemm_tc(&a, ep, e->stream);
    return launch_gemm_mma(&a, ep, e->stream);
}

// the kernel sequence of one decode step (captured into a graph by the callers)
int run_step(wf_engine* e, int B, int feed_back) {
    const wf_model_config& c = e->cfg;
    cudaStream_t s = e->stream;
    const long long launches0 = launch_count();
    int rc;
    if ((rc = launch_embed(e->embed, e->d_tok, B, e->hidden, e->hid, s))) return rc;
    const float attn_scale = 1.0f / sqrtf((float)e->D);
    for (int l = 0; l < c.layers; ++l) {
        LayerW& L = e->layers[l];
        if ((rc = launch_rmsnorm_quant(e->hid, L.g_in, c.eps, B, e->hidden, nullptr, e->q8, e->a_scale, e->rowsum, s, e->pdl))) return rc;
            {
            const int seg[3] = {e->H * e->D, (e->H + e->Hkv) * e->D, e->qkv_rows};
            if ((rc = linear(e, L.qkv, e->qkv_rows, e->hidden, B, L.scales + 0, 3, seg, 0, e->qkv, e->qkv_rows, nullptr))) return rc;
        }
        if ((rc = rope_append(e->qkv, B, e->H, e->Hkv, e->D, e->pool, c.layers, l, c.page_size, e->d_pos, e->d_block_table, e->max_blocks,
                              e->cos_tab, e->sin_tab, c.max_seq, e->qrot, s, e->pdl, e->kv_scales))) return rc;
            if ((rc = attn_decode(e->qrot, e->pool, c.layers, l, e->Hkv, c.page_size, e->d_block_table, e->max_blocks, e->d_len, B, e->H,
                              attn_scale, e->attn, WF_BF16, e->attn_seq_hint, e->ws_attn, e->ws_attn_bytes, s, e->pdl, e->kv_scales))) return rc;
            if ((rc = launch_rmsnorm_quant(e->attn, L.g_attn, c.eps, B, e->attn_dim, nullptr, e->q8, e->a_scale, e->rowsum, s, e->pdl))) return rc;
            {
            const int seg[1] = {e->hidden};
            if ((rc = linear(e, L.o, e->hidden, e->attn_dim, B, L.scales + 3, 1, seg, WF_EPI_RESIDUAL, e->hid, e->hidden, e->hid))) return rc;
        }
        if ((rc = launch_rmsnorm_quant(e->hid, L.g_post, c.eps, B, e->hidden, nullptr, e->q8, e->a_scale, e->rowsum, s, e->pdl))) return rc;
            {
            const int seg[2] = {e->ffn, 2 * e->ffn};   // unused by RELU2_MUL (scales picked per 16-row block)
            if ((rc = linear(e, L.gu, 2 * e->ffn, e->hidden, B, L.scales + 4, 2, seg, WF_EPI_RELU2_MUL, e->gu, e->ffn, nullptr))) return rc;
        }
        if ((rc = launch_rmsnorm_quant(e->gu, L.g_ffn, c.eps, B, e->ffn, nullptr, e->q8, e->a_scale, e->rowsum, s, e->pdl))) return rc;
            {
            const int seg[1] = {e->hidden};
            if ((rc = linear(e, L.down, e->hidden, e->ffn, B, L.scales + 6, 1, seg, WF_EPI_RESIDUAL, e->hid, e->hidden, e->hid))) return rc;
        }
    }
    // final norm (quant outputs unused) + tied LM head + greedy argmax
    if ((rc = launch_rmsnorm_quant(e->hid, e->final_norm, c.eps, B, e->hidden, e->normed, e->q8, e->a_scale, e->rowsum, s, false))) return rc;
    if ((rc = launch_lm_head_argmax(e->normed, e->embed, B, c.vocab, e->hidden, e->logits, e->d_next, e->ws_lm, s))) return rc;
    if ((rc = launch_advance(e->d_pos, e->d_len, e->d_tok, e->d_next, B, feed_back, s))) return rc;
    e->launches = (int)(launch_count() - launches0);
    return WF_OK;
}

// The step on the persistent layer-tail kernel (layer_tail.cu), B <= 8, single GPU:
//   embed-phase (input_layernorm + q|k|v of layer 0), then per layer { fused attention, tail }, final norm + LM head.
// tail(l) = o -> gate|up (+ relu^2*up) -> down -> q|k|v of layer l+1 in ONE launch: 2 launches per layer.
// Every linear output is written once by the CTA that owns its rows (no split-K), so no accumulator is ever zeroed.
bool tail_path_ok(const wf_engine* e, int B) {
    if (!e->tail || e->tp.size > 1) return false;
    int maxK = e->hidden;
    if (e->attn_full > maxK) maxK = e->attn_full;
    if (e->ffn_full > maxK) maxK = e->ffn_full;
    if (B > e->small_max) return e->tail_batch && layer_tail_batch_fits(B, maxK) && B <= sm_count();
    const int sms = sm_count();
    int maxrb = (e->qkv_rows / 16 + sms - 1) / sms;
    const int gu = ((e->ffn / 16 + sms - 1) / sms) * 2;
    if (gu > maxrb) maxrb = gu;
    return layer_tail_fits(B, maxK, maxrb);
}

int run_step_tail(wf_engine* e, int B, int feed_back) {
    const wf_model_config& c = e->cfg;
    cudaStream_t s = e->stream;
    const bool pdl = e->pdl;
    const long long launches0 = launch_count();
    const int MB = c.max_batch;
    float* asc[4] = {e->lin_asc, e->lin_asc + MB, e->lin_asc + 2 * MB, e->lin_asc + 3 * MB};
    int32_t* rs[4] = {e->lin_rs, e->lin_rs + MB, e->lin_rs + 2 * MB, e->lin_rs + 3 * MB};
    __nv_bfloat16 *hA = e->hid, *hB = e->hid2;
    int rc;
    const int fam = B <= e->small_max ? 0 : 1;
    auto base = [&](int site) {
        TailArgs t = {};
        t.M = B; t.bar = e->tail_bar[fam] + (size_t)site * LT_MAX_PHASES * (fam + 1); t.epoch = e->tail_epoch + fam; t.err = e->d_err;
        if (e->trace && site < c.layers) t.trace = e->trace + (size_t)32 * site;
        t.lam = e->tail_lam;
        return t;
    };
    auto launch_tail = [&](const TailArgs& t, bool use_pdl) { return fam ? launch_layer_tail_batch(t, s, use_pdl) : launch_layer_tail(t, s, use_pdl); };
    auto glue = [&](int mode, int K, const __nv_bfloat16* gamma, int slot) {
        GlueParams g = {};
        g.mode = mode; g.M = B; g.K = K; g.eps = c.eps; g.gamma = gamma; g.wait_flag = -1; g.tp = e->tp;
        g.q_out = fam ? e->q8c : e->q8; g.scale_out = asc[slot]; g.rowsum_out = rs[slot];      // batched form: the rows are published once, in global memory
        return g;
    };
    auto resid_from = [&](GlueParams& g, const __nv_bfloat16* in, __nv_bfloat16* out, const float* acc, int slot, const float* wscale) {
        g.hid_in = in; g.hid_out = out; g.acc = acc; g.acc_ld = e->hidden; g.acc_rowsum = rs[slot]; g.acc_ascale = asc[slot]; g.acc_wscale = wscale;
    };
    auto qkv_phase = [&](TailPhaseArgs& p, const LayerW& L) {
        p.w = L.qkv; p.N = e->qkv_rows; p.K = e->hidden; p.acc = e->acc_qkv; p.acc_ld = e->qkv_rows; p.lin_ascale = asc[0]; p.lin_rowsum = rs[0];
    };
    {   // embed -> input_layernorm -> q|k|v of layer 0
        TailArgs t = base(c.layers);
        t.nph = 1; t.bump_epoch = 1;
        GlueParams g = glue(GL_EMBED, e->hidden, e->layers[0].g_in, 0);
        g.embed = e->embed; g.tokens = e->d_tok; g.hid_out = hB;
        t.ph[0].g = g;
        qkv_phase(t.ph[0], e->layers[0]);
        if ((rc = launch_tail(t, false))) return rc;
    }
    const float attn_scale = 1.0f / sqrtf((float)e->D);
    for (int l = 0; l < c.layers; ++l) {
        LayerW& L = e->layers[l];
        {
            FusedAttnArgs a = {};
            a.pool = e->pool; a.kv_scales = e->kv_scales; a.layers = c.layers; a.layer = l; a.Hkv = e->Hkv; a.page_size = c.page_size;
            a.block_table = e->d_block_table; a.max_blocks = e->max_blocks; a.seq_lens = e->d_len; a.positions = e->d_pos;
            a.B = B; a.H = e->H; a.scale = attn_scale; a.out = e->attn;
            a.n_out = 1; a.out_ld = e->attn_full; a.out_col0 = 0; a.signal_flag = -1; a.tp = e->tp;
            a.qkv_acc = e->acc_qkv; a.qkv_ld = e->qkv_rows; a.qkv_rowsum = rs[0]; a.qkv_ascale = asc[0]; a.qkv_wscale = L.scales;
            a.cos_tab = e->cos_tab; a.sin_tab = e->sin_tab; a.max_pos = c.max_seq;
            a.merge_cnt = e->merge_cnt;
            a.max_seq_hint = e->attn_seq_hint; a.ws = e->ws_attn; a.ws_bytes = e->ws_attn_bytes;
            if (e->trace) a.trace = e->trace + (size_t)32 * (4 * c.layers + 4 + l);
            if ((rc = attn_decode_fused(a, s, pdl))) return rc;
        }
        TailArgs t = base(l);
        const bool last = (l + 1 == c.layers);
        t.nph = last ? 3 : 4;
        {   // o: attn_sub_norm over the attention output
            TailPhaseArgs& p = t.ph[0];
            p.g = glue(GL_PLAIN, e->attn_full, L.g_attn, 1);
            p.g.src = e->attn;
            p.w = L.o; p.N = e->hidden; p.K = e->attn_dim; p.acc = e->acc_o; p.acc_ld = e->hidden; p.lin_ascale = asc[1]; p.lin_rowsum = rs[1];
        }
        {   // gate|up: hidden + o -> post_attention_layernorm; epilogue relu^2(gate) * up -> h
            TailPhaseArgs& p = t.ph[1];
            p.g = glue(GL_RESID, e->hidden, L.g_post, 2);
            resid_from(p.g, hB, hA, e->acc_o, 1, L.scales + 3);
            p.w = L.gu; p.N = 2 * e->ffn; p.K = e->hidden; p.relu2 = true; p.lin_ascale = asc[2]; p.lin_rowsum = rs[2];
            p.h_out = e->h_full; p.h_ld = e->ffn_full; p.h_wscale = L.scales + 4;
        }
        {   // down: ffn_sub_norm over h
            TailPhaseArgs& p = t.ph[2];
            p.g = glue(GL_PLAIN, e->ffn_full, L.g_ffn, 3);
            p.g.src = e->h_full;
            p.w = L.down; p.N = e->hidden; p.K = e->ffn; p.acc = e->acc_down; p.acc_ld = e->hidden; p.lin_ascale = asc[3]; p.lin_rowsum = rs[3];
        }
        if (!last) {   // q|k|v of the next layer: hidden + down -> input_layernorm
            TailPhaseArgs& p = t.ph[3];
            p.g = glue(GL_RESID, e->hidden, e->layers[l + 1].g_in, 0);
            resid_from(p.g, hA, hB, e->acc_down, 3, L.scales + 6);
            qkv_phase(p, e->layers[l + 1]);
        }
        if ((rc = launch_tail(t, pdl))) return rc;
    }
    {   // hidden + down-proj -> final norm -> tied LM head + greedy argmax
        GlueParams g = glue(GL_RESID, e->hidden, e->final_norm, 0);
        resid_from(g, hA, hB, e->acc_down, 3, e->layers[c.layers - 1].scales + 6);
        g.y_out = e->normed; g.scale_out = nullptr; g.rowsum_out = nullptr;
        if ((rc = launch_glue_rows(g, s, pdl))) return rc;
    }
    TpLm tl = {};
    tl.tp = e->tp; tl.signal_flag = 4 * c.layers;
    if ((rc = launch_lm_head_argmax_tp(e->normed, e->embed, B, c.vocab, e->hidden, e->logits, c.vocab, 0, e->d_next, e->ws_lm, &tl, s))) return rc;
    if ((rc = launch_advance_tp(e->d_pos, e->d_len, e->d_tok, e->d_next, B, feed_back, e->d_step, &tl, s))) return rc;
    e->launches = (int)(launch_count() - launches0);
    return WF_OK;
}

// The same step as phase kernels (fused_common.cuh): every linear leaves its raw accumulator in an fp32 buffer and the
// NEXT phase applies scales / residual / relu^2*up / norm / re-quantisation in its prologue.
//   B <= 8 : 5 launches per layer (qkv, attention, o, gate|up, down), each linear = glue prologue + TMA-streamed GEMV
//   B  > 8 : 9 launches per layer (glue_rows + gemm_oneshot per linear, fused attention)
// Tensor parallel (tp.size = T > 1): q|k|v and gate|up are split by output rows (whole heads), attention and
// the KV pool by KV head, o and down by input columns.  The exchanges are fused into the producers (tp_comm.cuh):
//   B  attention stores its heads into attn_full of every rank            (all-gather, flag 4l+0)
//   C  o-proj REDs its exact partial sums into acc_o of every rank         (all-reduce, flag 4l+1)
//   D' relu^2(gate)*up of the local ffn slice goes to h_full of every rank (all-gather, flag 4l+2; one extra launch)
//   E  down-proj REDs into acc_down of every rank                          (all-reduce, flag 4l+3)
// and every rank normalises / quantises the gathered FULL vectors exactly like T = 1, so the step is bit-identical.
int run_step_fused(wf_engine* e, int B, int feed_back) {
    const wf_model_config& c = e->cfg;
    cudaStream_t s = e->stream;
    const bool pdl = e->pdl;
    const long long launches0 = launch_count();
    const int MB = c.max_batch;
    const int T = e->tp.size;
    const bool tp = T > 1;
    const bool hx = tp || e->hx;   // relu^2*up as its own small kernel (always under TP; measured faster than folding it into the down-proj glue at every batch size)
    float* asc[4] = {e->lin_asc, e->lin_asc + MB, e->lin_asc + 2 * MB, e->lin_asc + 3 * MB};
    int32_t* rs[4] = {e->lin_rs, e->lin_rs + MB, e->lin_rs + 2 * MB, e->lin_rs + 3 * MB};
    __nv_bfloat16 *hA = e->hid, *hB = e->hid2;
    int rc;
    int trace_slot = 0;
    // Row-parallel linears (o, down) under TP.  B <= 8: every CTA's prologue needs the whole reduced vector locally -> all-reduce
    // (RED into every rank).  B > 8: reduce-scatter (RED into the owner of the column only, 1/T of the NVLink packets) and the
    // per-row glue of the consumer pulls each column from its owner.
    const bool scatter = tp && B > e->small_max && (e->hidden % (8 * T)) == 0;
    const int own_cols = scatter ? e->hidden / T : 0;
    auto local_dest = [&](float* acc) { TpDest d = {}; d.n = 1; d.acc[0] = acc; d.signal_flag = -1; return d; };
    auto all_dest = [&](size_t off, int flag) {
        TpDest d = {};
        d.n = T; d.signal_flag = tp ? flag : -1; d.own_cols = own_cols;
        for (int r = 0; r < T; ++r) d.acc[r] = reinterpret_cast<float*>(e->tp.peer[r] + off);
        return d;
    };
    auto pull_from = [&](GlueParams& g, size_t off) {
        g.acc_own_cols = own_cols;
        if (scatter) for (int r = 0; r < T; ++r) g.acc_peer[r] = reinterpret_cast<const float*>(e->tp.peer[r] + off);
    };
    // one linear phase: prologue over the full activation vector of length Kfull, GEMV/GEMM over this rank's K columns
    auto phase = [&](GlueParams g, const uint8_t* w, int N, int Kfull, int K, int k_off, const TpDest& dest, int slot, bool first) -> int {
        g.M = B; g.K = Kfull; g.eps = c.eps; g.tp = e->tp;
        if (e->trace) { g.trace = e->trace + (size_t)trace_slot * 32; ++trace_slot; }
        if (B <= e->small_max && linear_small_fits(B, Kfull)) {
            SmallLinArgs a;
            a.g = g; a.w = w; a.N = N; a.K = K; a.k_off = k_off; a.dest = dest; a.acc_ld = N; a.lin_ascale = asc[slot]; a.lin_rowsum = rs[slot];
            return launch_linear_small(a, s, pdl && !first);
        }
        g.q_out = e->q8; g.scale_out = asc[slot]; g.rowsum_out = rs[slot];
        int r = launch_glue_rows(g, s, pdl && !first);
        if (r) return r;
        // a real dense contraction (33..64 tokens): tcgen05 int8 with split-K REDs; single-GPU engines only for now
        // tcgen05 split-K form (one owner per 128-row tile streams its whole K; the one-shot GEMM pays a RED round per 4 k-blocks): from 33
        // tokens up; from 9 tokens for K >= 8192 (70B shapes); from 24 tokens for >= 16 MB matrices.  Measured in profiles/r02q_summary.txt:
        // 70B bs16 32.2 -> 24.6 ms, 8B bs32 4.95 -> 4.84 ms, while 8B bs16 gate|up (29 MB, K = 4096) is faster on the one-shot GEMM.
        const size_t wbytes = (size_t)N * K / 4;
        // (17..24 rows: the two-pass form wins where the tcgen05 split-K form does not apply (K < 8192): 8B bs20/2K 4.19 -> 3.91 ms, bs24 4.35 -> 4.06;
        //  at the 70B widths it loses, 70B bs20/4K 20.0 -> 22.3: profiles/r02bc_summary.txt)
        const int stream_rows = e->decode_stream >= 2 ? 32 : (K < 8192 && e->decode_stream_max == 16 ? 24 : e->decode_stream_max);
        if (e->decode_stream && ((e->decode_stream_mask >> slot) & 1) && B <= stream_rows && gemm_stream_fits(B, N, K)) {
            OneshotArgs o;
            o.w = w; o.N = N; o.K = K; o.q = e->q8; o.M = B; o.ldq = Kfull; o.k_off = k_off; o.dest = dest; o.acc_ld = N; o.tp = e->tp;
            return launch_gemm_stream(o, s, pdl);
        }
        const bool use_tc = B >= e->decode_tc_min || (B >= 9 && K >= 8192) || (B >= 24 && wbytes >= ((size_t)16 << 20));
        // (tensor-parallel engines: the column-parallel linears — q|k|v, gate|up: local accumulator, no exchange — take it too; the
        //  row-parallel ones RED into peers and stay on the one-shot GEMM)
        const bool local_only = dest.n == 1 && dest.signal_flag < 0 && dest.own_cols == 0;
        if (e->decode_tc && local_only && use_tc && N % 8 == 0) return launch_gemm_tc_red(w, N, K, e->q8, B, Kfull, k_off, dest.acc[0], N, s, pdl);
        OneshotArgs o;
        o.w = w; o.N = N; o.K = K; o.q = e->q8; o.M = B; o.ldq = Kfull; o.k_off = k_off; o.dest = dest; o.acc_ld = N; o.tp = e->tp;
        return launch_gemm_oneshot(o, s, pdl);
    };
    const float attn_scale = 1.0f / sqrtf((float)e->D);
    for (int l = 0; l < c.layers; ++l) {
        LayerW& L = e->layers[l];
        const int fB = 4 * l, fC = 4 * l + 1, fD = 4 * l + 2, fE = 4 * l + 3;
        // ---- A: (embed | hidden + down-proj of the previous layer) -> input_layernorm -> qkv (this rank's heads)
        {
            GlueParams g = {};
            g.wait_flag = -1;
            if (l == 0) { g.mode = GL_EMBED; g.embed = e->embed; g.tokens = e->d_tok; }
            else {
                g.mode = GL_RESID; g.hid_in = hA; g.acc = e->acc_down; g.acc_ld = e->hidden;
                g.acc_rowsum = rs[3]; g.acc_ascale = asc[3]; g.acc_wscale = e->layers[l - 1].scales + 6;
                pull_from(g, e->off_acc_down);
                if (tp) g.wait_flag = fE - 4;
            }
            g.hid_out = hB; g.gamma = L.g_in;
            g.zero0 = e->acc_gu; g.zero0_n = B * 2 * e->ffn;
            if ((rc = phase(g, L.qkv, e->qkv_rows, e->hidden, e->hidden, 0, local_dest(e->acc_qkv), 0, l == 0))) return rc;
        }
        // ---- B: dequant + RoPE + KV append + paged attention + chunk merge (+ all-gather of the heads)
        {
            FusedAttnArgs a = {};
            a.pool = e->pool; a.kv_scales = e->kv_scales; a.layers = c.layers; a.layer = l; a.Hkv = e->Hkv; a.page_size = c.page_size;
            a.block_table = e->d_block_table; a.max_blocks = e->max_blocks; a.seq_lens = e->d_len; a.positions = e->d_pos;
            a.B = B; a.H = e->H; a.scale = attn_scale; a.out = e->attn;
            a.n_out = T; a.out_ld = e->attn_full; a.out_col0 = e->tp.rank * e->attn_dim; a.signal_flag = fB; a.tp = e->tp;
            for (int r = 0; r < T; ++r) a.out_peer[r] = e->tp.peer[r] + e->off_attn;
            a.qkv_acc = e->acc_qkv; a.qkv_ld = e->qkv_rows; a.qkv_rowsum = rs[0]; a.qkv_ascale = asc[0]; a.qkv_wscale = L.scales;
            a.cos_tab = e->cos_tab; a.sin_tab = e->sin_tab; a.max_pos = c.max_seq;
            a.merge_cnt = e->merge_cnt;
            // acc_down of the previous layer is cleared here — except under TP, where a peer may still be pulling / a slower
            // rank still reading it: there phase C clears it (it sits behind the attention all-gather flags of ALL ranks)
            a.zero0 = tp ? nullptr : e->acc_down; a.zero0_n = tp ? 0 : B * e->hidden;
            a.max_seq_hint = e->attn_seq_hint; a.ws = e->ws_attn; a.ws_bytes = e->ws_attn_bytes;
            if (e->trace) a.trace = e->trace + (size_t)32 * (4 * c.layers + 4 + l);
            if ((rc = attn_decode_fused(a, s, pdl))) return rc;
        }
        // ---- C: attn_sub_norm over the FULL attention vector -> o (this rank's input columns; RED = all-reduce)
        {
            GlueParams g = {};
            g.mode = GL_PLAIN; g.src = e->attn; g.gamma = L.g_attn; g.wait_flag = tp ? fB : -1;
            g.zero0 = e->acc_qkv; g.zero0_n = B * e->qkv_rows;
            if (tp) { g.zero1 = e->acc_down; g.zero1_n = B * e->hidden; }
            if ((rc = phase(g, L.o, e->hidden, e->attn_full, e->attn_dim, e->tp.rank * e->attn_dim, all_dest(e->off_acc_o, fC), 1, false))) return rc;
        }
        // ---- D: hidden + o-proj -> post_attention_layernorm -> gate|up (this rank's ffn rows)
        {
            GlueParams g = {};
            g.mode = GL_RESID; g.hid_in = hB; g.hid_out = hA; g.acc = e->acc_o; g.acc_ld = e->hidden;
            g.acc_rowsum = rs[1]; g.acc_ascale = asc[1]; g.acc_wscale = L.scales + 3; g.gamma = L.g_post; g.wait_flag = tp ? fC : -1;
            pull_from(g, e->off_acc_o);
            if ((rc = phase(g, L.gu, 2 * e->ffn, e->hidden, e->hidden, 0, local_dest(e->acc_gu), 2, false))) return rc;
        }
        // ---- D' (TP only): relu^2(gate)*up of the local slice -> h_full of every rank
        if (hx) {
            HExchangeArgs h = {};
            h.acc = e->acc_gu; h.M = B; h.ffn_local = e->ffn; h.ffn_full = e->ffn_full; h.col0 = e->tp.rank * e->ffn;
            h.rowsum = rs[2]; h.ascale = asc[2]; h.wscale = L.scales + 4;
            h.n_dst = T; h.signal_flag = tp ? fD : -1; h.tp = e->tp;
            for (int r = 0; r < T; ++r) h.dst[r] = reinterpret_cast<__nv_bfloat16*>(e->tp.peer[r] + e->off_h);
            if ((rc = launch_h_exchange(h, s, pdl))) return rc;
        }
        // ---- E: relu^2(gate)*up -> ffn_sub_norm over the FULL ffn vector -> down (RED = all-reduce)
        {
            GlueParams g = {};
            if (hx) { g.mode = GL_PLAIN; g.src = e->h_full; g.wait_flag = tp ? fD : -1; }
            else {
                g.mode = GL_RELU2; g.acc = e->acc_gu; g.acc_ld = 2 * e->ffn; g.wait_flag = -1;
                g.acc_rowsum = rs[2]; g.acc_ascale = asc[2]; g.acc_wscale = L.scales + 4;
            }
            g.gamma = L.g_ffn;
            g.zero0 = e->acc_o; g.zero0_n = B * e->hidden;
            if ((rc = phase(g, L.down, e->hidden, e->ffn_full, e->ffn, e->tp.rank * e->ffn, all_dest(e->off_acc_down, fE), 3, false))) return rc;
        }
    }
    // hidden + down-proj -> final norm -> tied LM head (this rank's vocab slice) + greedy argmax
    const int fLM = 4 * c.layers;
    {
        GlueParams g = {};
        g.mode = GL_RESID; g.M = B; g.K = e->hidden; g.eps = c.eps; g.tp = e->tp; g.wait_flag = tp ? fLM - 1 : -1;
        g.hid_in = hA; g.hid_out = hB; g.acc = e->acc_down; g.acc_ld = e->hidden;
        g.acc_rowsum = rs[3]; g.acc_ascale = asc[3]; g.acc_wscale = e->layers[c.layers - 1].scales + 6;
        pull_from(g, e->off_acc_down);
        g.gamma = e->final_norm; g.y_out = e->normed; g.q_out = e->q8;
        if ((rc = launch_glue_rows(g, s, pdl))) return rc;
    }
    TpLm tl = {};
    tl.tp = e->tp; tl.signal_flag = fLM;
    for (int r = 0; r < T; ++r) tl.cand_peer[r] = e->tp.peer[r] + e->off_cand;
    if ((rc = launch_lm_head_argmax_tp(e->normed, e->embed + (size_t)e->v_lo * e->hidden, B, e->v_n, e->hidden, e->logits + e->v_lo, c.vocab, e->v_lo,
                                       e->d_next, e->ws_lm, &tl, s))) return rc;
    if ((rc = launch_advance_tp(e->d_pos, e->d_len, e->d_tok, e->d_next, B, feed_back, e->d_step, &tl, s))) return rc;
    e->launches = (int)(launch_count() - launches0);
    return WF_OK;
}

// host-side bookkeeping before a step: make sure every active sequence has a page for its next token
int prepare_pages(wf_engine* e, int B) {
    const int ps = e->cfg.page_size;
    for (int b = 0; b < B; ++b) {
        const int pos = e->h_pos[b];
        if (pos >= e->cfg.max_seq) { set_error("sequence %d reached max_seq %d", b, e->cfg.max_seq); return WF_ERR_INVALID; }
        const int blk = pos / ps;
        if (e->h_block_table[(size_t)b * e->max_blocks + blk] < 0) {
            if (e->free_pages.empty()) { set_error("KV pool exhausted (%d pages)", e->cfg.max_pages); return WF_ERR_INVALID; }
            e->h_block_table[(size_t)b * e->max_blocks + blk] = e->free_pages.back();
            e->free_pages.pop_back();
            e->table_dirty = true;
        }
    }
    if (e->table_dirty) {
        // unused entries point at page 0 so that a stale lookup can never leave the pool
        std::vector<int32_t> tmp(e->h_block_table);
        for (auto& v : tmp) if (v < 0) v = 0;
        if (cuda_check(cudaMemcpyAsync(e->d_block_table, tmp.data(), tmp.size() * 4, cudaMemcpyHostToDevice, e->stream), "block table H2D")) return WF_ERR_CUDA;
        if (cuda_check(cudaStreamSynchronize(e->stream), "sync")) return WF_ERR_CUDA;   // tmp is pageable
        e->table_dirty = false;
    }
    return WF_OK;
}

int step_dispatch(wf_engine* e, int B, int feed_back) {
    if (!e->fused) return run_step(e, B, feed_back);
    if (tail_path_ok(e, B)) return run_step_tail(e, B, feed_back);
    return run_step_fused(e, B, feed_back);
}

int seq_hint_for(const wf_engine* e, int B) {
    int mx = 0;
    for (int b = 0; b < B; ++b) mx = e->h_pos[b] + 1 > mx ? e->h_pos[b] + 1 : mx;
    // round up to a power of two >= 256 so that captured graphs survive many steps
    int h = 256;
    while (h < mx) h <<= 1;
    if (h > e->cfg.max_seq) h = e->cfg.max_seq;
    return h;
}

int step_graph(wf_engine* e, int B, int feed_back, int use_graph) {
    const int hint = seq_hint_for(e, B);
    // The phase kernels keep a steady state between steps: every split-K accumulator is zero on entry because a LATER phase of the
    // previous step cleared it.  The tail kernels (plain stores by the owner of a row, nothing is ever cleared) do not maintain that
    // state, so a step that switches from them to the phase kernels (e.g. batch 1 -> batch 3 on one engine) clears the four
    // accumulators first.  (Found by test_batched_equals_single_sequences once batches 2..8 moved to the batched kernels.)
    {
        const int fam = !e->fused ? 0 : (tail_path_ok(e, B) ? 1 : 2);
        if (fam == 2 && e->last_family != 2) {   // (last_family: 0 per-op, 2 phase kernels, 16 + B tail kernels at batch B)
            const size_t MB = (size_t)e->cfg.max_batch;
            if (cuda_check(cudaMemsetAsync(e->acc_qkv, 0, MB * e->qkv_rows * 4, e->stream), "acc clear") ||
                cuda_check(cudaMemsetAsync(e->acc_gu, 0, MB * 2 * e->ffn * 4, e->stream), "acc clear") ||
                cuda_check(cudaMemsetAsync(e->acc_o, 0, MB * e->hidden * 4, e->stream), "acc clear") ||
                cuda_check(cudaMemsetAsync(e->acc_down, 0, MB * e->hidden * 4, e->stream), "acc clear")) return WF_ERR_CUDA;
        }
        // The batch-1 tail kernel's hand-off buffers (acc_o, acc_down, h) must hold the sentinel wherever their owner has not stored
        // this layer's data yet; in steady state the kernel re-arms them itself, on entry from any other step form they are filled here.
        const int famb = fam == 1 ? 16 + B : fam;
        if (fam == 1 && e->last_family != famb) {
            const size_t MB = (size_t)e->cfg.max_batch;
            if (cuda_check(cudaMemsetAsync(e->acc_o, 0xff, MB * e->hidden * 4, e->stream), "hand-off arm") ||
                cuda_check(cudaMemsetAsync(e->acc_down, 0xff, MB * e->hidden * 4, e->stream), "hand-off arm") ||
                cuda_check(cudaMemsetAsync(e->h_full, 0xff, MB * e->ffn_full * 2, e->stream), "hand-off arm")) return WF_ERR_CUDA;
        }
        e->last_family = famb;
    }
    if (!use_graph) {
        e->attn_seq_hint = hint;
        return step_dispatch(e, B, feed_back);
    }
    const int gi = feed_back ? 1 : 0;
    if (!e->graph_exec[gi] || e->graph_B[gi] != B || e->graph_hint[gi] != hint) {
        if (e->graph_exec[gi]) { cudaGraphExecDestroy(e->graph_exec[gi]); e->graph_exec[gi] = nullptr; }
        e->attn_seq_hint = hint;
        // one eager run first: sets function attributes (not allowed during capture) — on scratch state
        // this would advance positions, so capture directly but pre-set attributes through a dry launch
        // with B rows on the real buffers is unsafe; instead we rely on launchers being attribute-idempotent
        // and run them once outside capture in wf_engine_finalize_weights (warm-up step on slot state 0).
        cudaGraph_t graph = nullptr;
        if (cuda_check(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal), "begin capture")) return WF_ERR_CUDA;
        int rc = step_dispatch(e, B, feed_back);
        cudaError_t ce = cudaStreamEndCapture(e->stream, &graph);
        if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (cuda_check(ce, "end capture")) return WF_ERR_CUDA;
        ce = cudaGraphInstantiate(&e->graph_exec[gi], graph, 0);
        cudaGraphDestroy(graph);
        if (cuda_check(ce, "graph instantiate")) return WF_ERR_CUDA;
        e->graph_B[gi] = B;
        e->graph_hint[gi] = hint;
    }
    return cuda_check(cudaGraphLaunch(e->graph_exec[gi], e->stream), "graph launch");
}

}  // namespace

extern "C" int wf_engine_create(const wf_model_config* cfg, wf_engine** out) {
    if (!cfg || !out) { set_error("wf_engine_create: null argument"); return WF_ERR_INVALID; }
    const wf_model_config& c = *cfg;
    if (c.head_dim != 128) { set_error("engine: head_dim must be 128"); return WF_ERR_UNSUPPORTED; }
    const int T = c.tp_size <= 0 ? 1 : c.tp_size;
    if (T > TP_MAX || c.tp_rank < 0 || c.tp_rank >= T) { set_error("engine: tp_size must be 1..%d and 0 <= tp_rank < tp_size", TP_MAX); return WF_ERR_INVALID; }
    if (c.heads % T || c.kv_heads % T || c.ffn % (16 * T) || (c.ffn / T) % 128) {
        set_error("engine: heads (%d), kv_heads (%d) and ffn/128 (%d) must be divisible by tp_size %d (SURVEY §8e: the 2B shape only shards over 1 or 5 GPUs)",
                  c.heads, c.kv_heads, c.ffn / 128, T);
        return WF_ERR_UNSUPPORTED;
    }
    if (c.heads % c.kv_heads || c.heads / c.kv_heads > 8) { set_error("engine: heads/kv_heads must be an integer <= 8"); return WF_ERR_UNSUPPORTED; }
    if (c.hidden % 128 || c.ffn % 128 || (c.heads * c.head_dim) % 128) { set_error("engine: hidden, ffn, heads*head_dim must be multiples of 128"); return WF_ERR_UNSUPPORTED; }
    if (c.ffn % 16 || c.page_size % 16 || c.page_size <= 0) { set_error("engine: page_size must be a positive multiple of 16"); return WF_ERR_UNSUPPORTED; }
    if (c.max_batch < 1 || c.max_batch > 64) { set_error("engine: max_batch must be 1..64"); return WF_ERR_UNSUPPORTED; }
    if (c.max_seq % c.page_size) { set_error("engine: max_seq must be a multiple of page_size"); return WF_ERR_INVALID; }
    wf_engine* e = new wf_engine();
    e->cfg = c;
    e->cfg.tp_size = T;
    e->H = c.heads / T; e->Hkv = c.kv_heads / T; e->D = c.head_dim; e->hidden = c.hidden; e->ffn = c.ffn / T;
    e->qkv_rows = (e->H + 2 * e->Hkv) * e->D;
    e->attn_dim = e->H * e->D;
    e->attn_full = c.heads * c.head_dim;
    e->ffn_full = c.ffn;
    memset(&e->tp, 0, sizeof e->tp);
    e->tp.rank = c.tp_rank; e->tp.size = T;
    {   // vocab slice of the tied LM head: whole 128-row CTAs per rank
        const int slice = (((c.vocab + T - 1) / T) + 127) / 128 * 128;
        e->v_lo = c.tp_rank * slice;
        if (e->v_lo > c.vocab) e->v_lo = c.vocab;
        e->v_n = c.vocab - e->v_lo < slice ? c.vocab - e->v_lo : slice;
        if (e->v_n <= 0) { delete e; set_error("engine: vocab %d too small for tp_size %d", c.vocab, T); return WF_ERR_UNSUPPORTED; }
    }
    const char* env_pdl = getenv("WF_PDL");
    e->pdl = env_pdl && atoi(env_pdl) != 0;
    const char* env_algo = getenv("WF_LINEAR_ALGO");
    e->linear_algo = env_algo ? atoi(env_algo) : 0;
    const char* env_fused = getenv("WF_FUSED");
    e->fused = env_fused ? atoi(env_fused) != 0 : 1;
    if (T > 1) e->fused = 1;                   // the exchanges live in the phase kernels
    { const char* x = getenv("WF_HX"); if (x) e->hx = atoi(x) != 0; }
    { const char* x = getenv("WF_TAIL"); if (x) e->tail = atoi(x) != 0; }
    { const char* x = getenv("WF_TAIL_LAM"); if (x) e->tail_lam = atoi(x); }
    { const char* x = getenv("WF_SMALL_MAX"); if (x) { e->small_max = atoi(x); if (e->small_max < 0) e->small_max = 0; if (e->small_max > 8) e->small_max = 8; } }
    { const char* x = getenv("WF_TAIL_BATCH"); if (x) e->tail_batch = atoi(x) != 0; }
    { const char* x = getenv("WF_DECODE_TC"); if (x) e->decode_tc = atoi(x) != 0; }
    { const char* x = getenv("WF_DECODE_TC_MIN"); if (x) e->decode_tc_min = atoi(x) < 2 ? 2 : atoi(x); }
    { const char* x = getenv("WF_DECODE_STREAM"); if (x) e->decode_stream = atoi(x); }
    { const char* x = getenv("WF_DECODE_STREAM_MASK"); if (x) e->decode_stream_mask = atoi(x) & 15; }
    { const char* x = getenv("WF_DECODE_STREAM_MAX"); if (x) e->decode_stream_max = atoi(x) < 1 ? 1 : (atoi(x) > 32 ? 32 : atoi(x)); }
    if (e->fused && !env_pdl) e->pdl = true;   // the phase kernels are built around programmatic dependent launch
    // Tensor-parallel engines keep the one-shot / tcgen05 linears: with the streaming GEMM under the exchanges TP = 2 at the 8B layer shape, batch 16, was
    // bit-identical to TP = 1 in 12 of 14 runs only (profiles/r02bf_summary.txt; 0 mismatches in 33 runs without it, 0 in 10 x 7 single-GPU runs) — a
    // rare ordering problem between its REDs into peer accumulators and the flag protocol that was not tracked down.  WF_DECODE_STREAM=1 forces it.
    if (T > 1 && !getenv("WF_DECODE_STREAM")) e->decode_stream = 0;
    // Batch 1 on shapes whose ffn row is too long for the hand-off form of the tail kernel (register glue: <= 9216 elements; the 8B / 70B widths)
    // goes through the row glue + streaming batched GEMM instead of the grid-barrier tail kernel: 8B bs1 1.97 -> 1.86 ms, 70B bs1 9.32 -> 8.22 ms
    // (profiles/r02az_summary.txt).  Single-GPU engines only: under tensor parallelism batch 1 keeps the all-RED exchange of the GEMV-style kernels.
    if (!getenv("WF_SMALL_MAX") && T == 1 && e->decode_stream) {
        int maxK = e->hidden > e->attn_full ? e->hidden : e->attn_full;
        if (e->ffn_full > maxK) maxK = e->ffn_full;
        if (maxK > 9216) e->small_max = 0;
    }
    int rc = WF_OK;
    auto fail = [&](int r) { wf_engine_destroy(e); return r; };
    if (cuda_check(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking), "stream create")) return fail(WF_ERR_CUDA);
    init_all_kernel_attrs();
    fused_init_attrs();
    cudaGetLastError();
    const int B = c.max_batch;
    e->layers.resize(c.layers);
    for (int l = 0; l < c.layers; ++l) {
        LayerW& L = e->layers[l];
        if ((rc = dalloc(e, &L.qkv, wf_tile_bytes(e->qkv_rows, e->hidden)))) return fail(rc);
        if ((rc = dalloc(e, &L.o, wf_tile_bytes(e->hidden, e->attn_dim)))) return fail(rc);
        if ((rc = dalloc(e, &L.gu, wf_tile_bytes(2 * e->ffn, e->hidden)))) return fail(rc);
        if ((rc = dalloc(e, &L.down, wf_tile_bytes(e->hidden, e->ffn)))) return fail(rc);
        if ((rc = dalloc(e, &L.scales, 8))) return fail(rc);
        if ((rc = dalloc(e, &L.g_in, e->hidden))) return fail(rc);
        if ((rc = dalloc(e, &L.g_post, e->hidden))) return fail(rc);
        if ((rc = dalloc(e, &L.g_attn, e->attn_full))) return fail(rc);
        if ((rc = dalloc(e, &L.g_ffn, e->ffn_full))) return fail(rc);
    }
    if ((rc = dalloc(e, &e->embed, (size_t)c.vocab * c.hidden))) return fail(rc);
    if ((rc = dalloc(e, &e->final_norm, c.hidden))) return fail(rc);
    if ((rc = dalloc(e, &e->cos_tab, (size_t)c.max_seq * (c.head_dim / 2)))) return fail(rc);
    if ((rc = dalloc(e, &e->sin_tab, (size_t)c.max_seq * (c.head_dim / 2)))) return fail(rc);
    // rope tables, same construction as RotaryEmbedding.__init__ (direct_inference_full.py:31-45):
    // fp32 inv_freq, fp32 outer product, cos/sin, round to bf16.  (The host layer may overwrite them with
    // torch-computed tables through wf_engine_set_tensor("rope_cos"/"rope_sin") for bit-level parity.)
    {
        const int half = c.head_dim / 2;
        std::vector<__nv_bfloat16> hc((size_t)c.max_seq * half), hs((size_t)c.max_seq * half);
        for (int i = 0; i < half; ++i) {
            const float inv = 1.0f / powf(c.rope_theta, (float)(2 * i) / (float)c.head_dim);
            for (int p = 0; p < c.max_seq; ++p) {
                const float f = (float)p * inv;
                hc[(size_t)p * half + i] = __float2bfloat16_rn(cosf(f));
                hs[(size_t)p * half + i] = __float2bfloat16_rn(sinf(f));
            }
        }
        if (cuda_check(cudaMemcpy(e->cos_tab, hc.data(), hc.size() * 2, cudaMemcpyHostToDevice), "rope cos")) return fail(WF_ERR_CUDA);
        if (cuda_check(cudaMemcpy(e->sin_tab, hs.data(), hs.size() * 2, cudaMemcpyHostToDevice), "rope sin")) return fail(WF_ERR_CUDA);
    }
    // KV pool
    const size_t pool_elems = (size_t)c.max_pages * c.layers * 2 * e->Hkv * c.page_size * c.head_dim;
    if (c.kv_dtype == WF_I8) {   // int8 codes + one fp32 scale per (token, kv head), one allocation (the wf_kv_layout convention)
        if (c.head_dim != 128) { set_error("int8 KV needs head_dim 128"); return fail(WF_ERR_UNSUPPORTED); }
        unsigned char* raw = nullptr;
        const size_t bytes = pool_elems + pool_elems / c.head_dim * sizeof(float);
        if ((rc = dalloc(e, &raw, bytes))) return fail(rc);
        if (cuda_check(cudaMemset(raw, 0, bytes), "pool memset")) return fail(WF_ERR_CUDA);
        e->pool = reinterpret_cast<__nv_bfloat16*>(raw);
        e->kv_scales = reinterpret_cast<float*>(raw + pool_elems);
    } else {
        if (c.kv_dtype != 0 && c.kv_dtype != WF_BF16) { set_error("kv_dtype must be WF_BF16 or WF_I8"); return fail(WF_ERR_INVALID); }
        if ((rc = dalloc(e, &e->pool, pool_elems))) return fail(rc);
        if (cuda_check(cudaMemset(e->pool, 0, pool_elems * 2), "pool memset")) return fail(WF_ERR_CUDA);
    }
    e->max_blocks = c.max_seq / c.page_size;
    if ((rc = dalloc(e, &e->d_block_table, (size_t)B * e->max_blocks))) return fail(rc);
    if ((rc = dalloc(e, &e->d_pos, B)) || (rc = dalloc(e, &e->d_len, B)) || (rc = dalloc(e, &e->d_tok, B)) || (rc = dalloc(e, &e->d_next, B))) return fail(rc);
    if (cuda_check(cudaMallocHost(&e->h_tok_pinned, B * 4), "pinned") || cuda_check(cudaMallocHost(&e->h_next_pinned, B * 4), "pinned") ||
        cuda_check(cudaMallocHost(&e->h_pos_pinned, 2 * B * 4), "pinned")) return fail(WF_ERR_CUDA);
    // activations
    int maxK = e->hidden;
    if (e->attn_full > maxK) maxK = e->attn_full;
    if (e->ffn_full > maxK) maxK = e->ffn_full;
    // exchange buffer: everything a peer writes into (tp_comm.cuh).  One cudaMalloc so that one IPC handle exports it.
    {
        auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
        size_t off = 0;
        e->off_attn = off; off = up(off + (size_t)B * e->attn_full * 2);
        e->off_h = off; off = up(off + (size_t)B * e->ffn_full * 2);
        e->off_acc_o = off; off = up(off + (size_t)B * e->hidden * 4);
        e->off_acc_down = off; off = up(off + (size_t)B * e->hidden * 4);
        e->off_cand = off; off = up(off + (size_t)B * TP_MAX * 8);
        e->off_flags = off; off = up(off + (size_t)(4 * c.layers + 1) * TP_MAX * 4);
        e->comm_bytes = off;
        if ((rc = dalloc(e, &e->comm, e->comm_bytes))) return fail(rc);
        if (cuda_check(cudaMemset(e->comm, 0, e->comm_bytes), "comm memset")) return fail(WF_ERR_CUDA);
        e->attn = reinterpret_cast<__nv_bfloat16*>(e->comm + e->off_attn);
        e->h_full = reinterpret_cast<__nv_bfloat16*>(e->comm + e->off_h);
        e->acc_o = reinterpret_cast<float*>(e->comm + e->off_acc_o);
        e->acc_down = reinterpret_cast<float*>(e->comm + e->off_acc_down);
        e->tp.peer[c.tp_rank] = e->comm;
        e->tp.off_flags = e->off_flags;
        int* cnt = nullptr;
        if ((rc = dalloc(e, &cnt, (size_t)4 * c.layers + 1)) || (rc = dalloc(e, &e->d_step, 1)) || (rc = dalloc(e, &e->d_err, 1))) return fail(rc);
        if (cuda_check(cudaMemset(cnt, 0, ((size_t)4 * c.layers + 1) * 4), "cnt memset") || cuda_check(cudaMemset(e->d_step, 0, 4), "step memset") ||
            cuda_check(cudaMemset(e->d_err, 0, 4), "err memset")) return fail(WF_ERR_CUDA);
        if (cuda_check(cudaMallocHost(&e->h_err_pinned, 4), "pinned")) return fail(WF_ERR_CUDA);
        *e->h_err_pinned = 0;
        e->tp.done_cnt = cnt;
        e->tp.step = e->d_step;
        e->tp.err = e->d_err;
        e->tp_attached = (T == 1);
    }
    if ((rc = dalloc(e, &e->hid, (size_t)B * e->hidden)) || (rc = dalloc(e, &e->normed, (size_t)B * e->hidden)) ||
        (rc = dalloc(e, &e->qkv, (size_t)B * e->qkv_rows)) || (rc = dalloc(e, &e->qrot, (size_t)B * e->attn_dim)) ||
        (rc = dalloc(e, &e->gu, (size_t)B * e->ffn)) ||
        (rc = dalloc(e, &e->q8, (size_t)B * maxK)) || (rc = dalloc(e, &e->q8c, (size_t)((maxK + 511) / 512) * QCHUNK_BYTES)) || (rc = dalloc(e, &e->a_scale, B)) || (rc = dalloc(e, &e->rowsum, B)) ||
        (rc = dalloc(e, &e->logits, (size_t)B * c.vocab))) return fail(rc);
    if ((rc = dalloc(e, &e->acc_qkv, (size_t)B * e->qkv_rows)) ||
        (rc = dalloc(e, &e->acc_gu, (size_t)B * 2 * e->ffn)) ||
        (rc = dalloc(e, &e->lin_asc, (size_t)4 * B)) || (rc = dalloc(e, &e->lin_rs, (size_t)4 * B)) ||
        (rc = dalloc(e, &e->hid2, (size_t)B * e->hidden)) || (rc = dalloc(e, &e->merge_cnt, (size_t)B * e->Hkv))) return fail(rc);
    if ((rc = dalloc(e, &e->tail_bar[0], (size_t)(c.layers + 1) * LT_MAX_PHASES)) || (rc = dalloc(e, &e->tail_bar[1], (size_t)(c.layers + 1) * 2 * LT_MAX_PHASES)) ||
        (rc = dalloc(e, &e->tail_epoch, 2))) return fail(rc);
    if (cuda_check(cudaMemset(e->tail_bar[0], 0, (size_t)(c.layers + 1) * LT_MAX_PHASES * 8), "bar memset") ||
        cuda_check(cudaMemset(e->tail_bar[1], 0, (size_t)(c.layers + 1) * 2 * LT_MAX_PHASES * 8), "bar memset") ||
        cuda_check(cudaMemset(e->tail_epoch, 0, 8), "epoch memset")) return fail(WF_ERR_CUDA);
    if (cuda_check(cudaMemset(e->acc_qkv, 0, (size_t)B * e->qkv_rows * 4), "acc memset") ||
        cuda_check(cudaMemset(e->acc_gu, 0, (size_t)B * 2 * e->ffn * 4), "acc memset") ||
        cuda_check(cudaMemset(e->merge_cnt, 0, (size_t)B * e->Hkv * 4), "cnt memset")) return fail(WF_ERR_CUDA);
    const int maxN = 2 * e->ffn > e->qkv_rows ? 2 * e->ffn : e->qkv_rows;
    e->ws_lin_bytes = wf_linear_workspace_bytes(B, maxN, maxK) + (size_t)maxN * 4;
    e->ws_attn_bytes = wf_attn_workspace_bytes(B, e->H, e->D, c.max_seq);
    e->ws_lm_bytes = wf_lm_head_workspace_bytes(B, c.vocab);
    {
        char *a = nullptr, *b2 = nullptr, *c2 = nullptr;
        if ((rc = dalloc(e, &a, e->ws_lin_bytes)) || (rc = dalloc(e, &b2, e->ws_attn_bytes)) || (rc = dalloc(e, &c2, e->ws_lm_bytes))) return fail(rc);
        e->ws_lin = a; e->ws_attn = b2; e->ws_lm = c2;
        if (cuda_check(cudaMemset(e->ws_lin, 0, e->ws_lin_bytes), "ws memset")) return fail(WF_ERR_CUDA);
    }
    if (getenv("WF_TRACE") && atoi(getenv("WF_TRACE"))) {
        if ((rc = dalloc(e, &e->trace, (size_t)32 * (5 * c.layers + 8)))) return fail(rc);
        cudaMemset(e->trace, 0, (size_t)32 * (5 * c.layers + 8) * 8);
    }
    e->h_block_table.assign((size_t)B * e->max_blocks, -1);
    e->h_pos.assign(B, 0);
    wf_engine_reset(e);
    *out = e;
    return WF_OK;
}

extern "C" void wf_engine_destroy(wf_engine* e) {
    if (!e) return;
    cudaDeviceSynchronize();
    invalidate_graphs(e);
    for (int r = 0; r < TP_MAX; ++r)
        if (e->peer_opened[r]) cudaIpcCloseMemHandle(e->peer_opened[r]);
    for (void* p : e->allocs) cudaFree(p);
    for (auto& L : e->layers)
        for (int i = 0; i < 7; ++i)
            if (L.stage[i]) cudaFree(L.stage[i]);
    if (e->h_tok_pinned) cudaFreeHost(e->h_tok_pinned);
    if (e->h_next_pinned) cudaFreeHost(e->h_next_pinned);
    if (e->h_pos_pinned) cudaFreeHost(e->h_pos_pinned);
    if (e->h_err_pinned) cudaFreeHost(e->h_err_pinned);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

extern "C" size_t wf_engine_device_bytes(const wf_engine* e) { return e ? e->dev_bytes : 0; }
extern "C" void* wf_engine_stream(wf_engine* e) { return e ? (void*)e->stream : nullptr; }
extern "C" int wf_engine_launches_per_step(const wf_engine* e) { return e ? e->launches : 0; }
// device-side spins are bounded (tp_comm.cuh, layer_tail.cu): a wait that gave up leaves a code in the error word
static int check_device_error(wf_engine* e) {
    if (!e->d_err) return WF_OK;
    if (cuda_check(cudaMemcpyAsync(e->h_err_pinned, e->d_err, 4, cudaMemcpyDeviceToHost, e->stream), "err D2H")) return WF_ERR_CUDA;
    if (cuda_check(cudaStreamSynchronize(e->stream), "engine sync")) return WF_ERR_CUDA;
    if (*e->h_err_pinned) {
        const int code = *e->h_err_pinned;
        set_error("device-side wait timed out (code %d: %s) — a tensor-parallel peer died, the ranks' step counters disagree, or a grid barrier lost a CTA; "
                  "the engine state is undefined, destroy it", code, code >= 1000 ? "TP exchange flag" : "grid barrier");
        return WF_ERR_CUDA;
    }
    return WF_OK;
}
extern "C" int wf_engine_sync(wf_engine* e) { return e ? check_device_error(e) : WF_ERR_INVALID; }

extern "C" int wf_engine_set_linear(wf_engine* e, int layer, const char* name, const uint8_t* i2s, float weight_scale, int is_host) {
    if (!e || !name || !i2s) { set_error("wf_engine_set_linear: null argument"); return WF_ERR_INVALID; }
    if (layer < 0 || layer >= e->cfg.layers) { set_error("wf_engine_set_linear: layer %d out of range", layer); return WF_ERR_INVALID; }
    const int idx = lin_index(name);
    if (idx < 0) { set_error("wf_engine_set_linear: unknown projection '%s'", name); return WF_ERR_INVALID; }
    int N, K;
    lin_shape(e, idx, &N, &K);
    LayerW& L = e->layers[layer];
    const size_t bytes = (size_t)N * K / 4;
    if (!L.stage[idx] && cuda_check(cudaMalloc((void**)&L.stage[idx], bytes), "stage alloc")) return WF_ERR_CUDA;
    if (cuda_check(cudaMemcpy(L.stage[idx], i2s, bytes, is_host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice), "stage copy")) return WF_ERR_CUDA;
    L.h_scales[idx] = weight_scale;
    e->finalized = false;
    return WF_OK;
}

extern "C" int wf_engine_set_tensor(wf_engine* e, int layer, const char* name, const void* data, size_t count, int is_host) {
    if (!e || !name || !data) { set_error("wf_engine_set_tensor: null argument"); return WF_ERR_INVALID; }
    __nv_bfloat16* dst = nullptr;
    size_t expect = 0;
    const std::string n(name);
    if (layer < 0) {
        if (n == "embed") { dst = e->embed; expect = (size_t)e->cfg.vocab * e->hidden; }
        else if (n == "final_norm") { dst = e->final_norm; expect = e->hidden; }
        else if (n == "rope_cos") { dst = e->cos_tab; expect = (size_t)e->cfg.max_seq * (e->D / 2); }
        else if (n == "rope_sin") { dst = e->sin_tab; expect = (size_t)e->cfg.max_seq * (e->D / 2); }
    } else if (layer < e->cfg.layers) {
        LayerW& L = e->layers[layer];
        if (n == "input_layernorm") { dst = L.g_in; expect = e->hidden; }
        else if (n == "post_attention_layernorm") { dst = L.g_post; expect = e->hidden; }
        else if (n == "attn_sub_norm") { dst = L.g_attn; expect = e->attn_full; }
        else if (n == "ffn_sub_norm") { dst = L.g_ffn; expect = e->ffn_full; }
    }
    if (!dst) { set_error("wf_engine_set_tensor: unknown tensor '%s' (layer %d)", name, layer); return WF_ERR_INVALID; }
    if (count != expect) { set_error("wf_engine_set_tensor: '%s' expects %zu elements, got %zu", name, expect, count); return WF_ERR_INVALID; }
    return cuda_check(cudaMemcpy(dst, data, count * 2, is_host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice), "set_tensor copy");
}

extern "C" int wf_engine_finalize_weights(wf_engine* e) {
    if (!e) return WF_ERR_INVALID;
    cudaStream_t s = e->stream;
    for (int l = 0; l < e->cfg.layers; ++l) {
        LayerW& L = e->layers[l];
        bool any = false, all = true;
        for (int i = 0; i < 7; ++i) { any |= L.stage[i] != nullptr; all &= L.stage[i] != nullptr; }
        if (!any) continue;   // synthetic or already finalized
        if (!all) { set_error("layer %d: all seven projections must be set before finalize", l); return WF_ERR_INVALID; }
        int rc;
        const int K4 = e->hidden / 4;
        // q|k|v: rows concatenated -> one I2_S matrix -> tiles
        {
            uint8_t* cat = nullptr;
            const size_t bq = (size_t)e->H * e->D * K4, bk = (size_t)e->Hkv * e->D * K4;
            if (cuda_check(cudaMalloc((void**)&cat, bq + 2 * bk), "cat alloc")) return WF_ERR_CUDA;
            cudaMemcpyAsync(cat, L.stage[0], bq, cudaMemcpyDeviceToDevice, s);
            cudaMemcpyAsync(cat + bq, L.stage[1], bk, cudaMemcpyDeviceToDevice, s);
            cudaMemcpyAsync(cat + bq + bk, L.stage[2], bk, cudaMemcpyDeviceToDevice, s);
            rc = wf_repack_i2s_to_tile(cat, e->qkv_rows, e->hidden, L.qkv, s);
            cudaStreamSynchronize(s);
            cudaFree(cat);
            if (rc) return rc;
        }
        if ((rc = wf_repack_i2s_to_tile(L.stage[3], e->hidden, e->attn_dim, L.o, s))) return rc;
        {
            uint8_t* il = nullptr;
            if (cuda_check(cudaMalloc((void**)&il, (size_t)2 * e->ffn * K4), "interleave alloc")) return WF_ERR_CUDA;
            rc = wf_interleave_rows16(L.stage[4], L.stage[5], e->ffn, K4, il, s);
            if (!rc) rc = wf_repack_i2s_to_tile(il, 2 * e->ffn, e->hidden, L.gu, s);
            cudaStreamSynchronize(s);
            cudaFree(il);
            if (rc) return rc;
        }
        if ((rc = wf_repack_i2s_to_tile(L.stage[6], e->hidden, e->ffn, L.down, s))) return rc;
        if (cuda_check(cudaMemcpyAsync(L.scales, L.h_scales, 7 * 4, cudaMemcpyHostToDevice, s), "scales")) return WF_ERR_CUDA;
        cudaStreamSynchronize(s);
        for (int i = 0; i < 7; ++i) { cudaFree(L.stage[i]); L.stage[i] = nullptr; }
    }
    e->finalized = true;
    invalidate_graphs(e);
    return cuda_check(cudaStreamSynchronize(s), "finalize");
}

extern "C" int wf_engine_init_synthetic(wf_engine* e, uint64_t seed) {
    if (!e) return WF_ERR_INVALID;
    cudaStream_t s = e->stream;
    int rc;
    for (int l = 0; l < e->cfg.layers; ++l) {
        LayerW& L = e->layers[l];
        const uint64_t sd = seed * 1000003ull + (uint64_t)l * 7919ull;
        const uint64_t sr = sd + (uint64_t)e->tp.rank * 0x9E3779B97F4A7C15ull;   // sharded matrices differ per rank, replicated tensors do not
        if ((rc = launch_fill_ternary(L.qkv, wf_tile_bytes(e->qkv_rows, e->hidden), sr + 1, s))) return rc;
        if ((rc = launch_fill_ternary(L.o, wf_tile_bytes(e->hidden, e->attn_dim), sr + 2, s))) return rc;
        if ((rc = launch_fill_ternary(L.gu, wf_tile_bytes(2 * e->ffn, e->hidden), sr + 3, s))) return rc;
        if ((rc = launch_fill_ternary(L.down, wf_tile_bytes(e->hidden, e->ffn), sr + 4, s))) return rc;
        // weight_scale ~ U(0.5, 2.5); deterministic LCG on the host
        uint64_t st = sd ^ 0xD1B54A32D192ED03ull;
        for (int i = 0; i < 7; ++i) {
            st = st * 6364136223846793005ull + 1442695040888963407ull;
            L.h_scales[i] = 0.5f + 2.0f * (float)((st >> 40) & 0xffff) / 65536.0f;
        }
        if (cuda_check(cudaMemcpyAsync(L.scales, L.h_scales, 7 * 4, cudaMemcpyHostToDevice, s), "scales")) return WF_ERR_CUDA;
        cudaStreamSynchronize(s);
        if ((rc = launch_fill_bf16(L.g_in, e->hidden, sd + 5, 1.0f, 0.1f, s))) return rc;
        if ((rc = launch_fill_bf16(L.g_post, e->hidden, sd + 6, 1.0f, 0.1f, s))) return rc;
        if ((rc = launch_fill_bf16(L.g_attn, e->attn_full, sd + 7, 1.0f, 0.1f, s))) return rc;
        if ((rc = launch_fill_bf16(L.g_ffn, e->ffn_full, sd + 8, 1.0f, 0.1f, s))) return rc;
    }
    if ((rc = launch_fill_bf16(e->embed, (size_t)e->cfg.vocab * e->hidden, seed + 11, 0.0f, 0.02f, s))) return rc;
    if ((rc = launch_fill_bf16(e->final_norm, e->hidden, seed + 12, 1.0f, 0.1f, s))) return rc;
    e->finalized = true;
    invalidate_graphs(e);
    return cuda_check(cudaStreamSynchronize(s), "init_synthetic");
}

extern "C" int wf_engine_reset(wf_engine* e) {
    if (!e) return WF_ERR_INVALID;
    const int B = e->cfg.max_batch;
    e->free_pages.clear();
    for (int p = e->cfg.max_pages - 1; p >= 0; --p) e->free_pages.push_back(p);   // pop_back hands out 0,1,2,... like the reference free list
    std::fill(e->h_block_table.begin(), e->h_block_table.end(), -1);
    std::fill(e->h_pos.begin(), e->h_pos.end(), 0);
    e->table_dirty = true;
    if (cuda_check(cudaMemsetAsync(e->d_pos, 0, B * 4, e->stream), "reset")) return WF_ERR_CUDA;
    std::vector<int32_t> ones(B, 1);
    if (cuda_check(cudaMemcpyAsync(e->d_len, ones.data(), B * 4, cudaMemcpyHostToDevice, e->stream), "reset")) return WF_ERR_CUDA;
    return cuda_check(cudaStreamSynchronize(e->stream), "reset");
}

extern "C" int wf_engine_set_context(wf_engine* e, int slot, int len, int fill_random) {
    if (!e || slot < 0 || slot >= e->cfg.max_batch || len < 0 || len >= e->cfg.max_seq) { set_error("wf_engine_set_context: bad slot/len"); return WF_ERR_INVALID; }
    const int ps = e->cfg.page_size;
    // release pages beyond the new length, allocate up to it
    for (int blk = 0; blk < e->max_blocks; ++blk) {
        int32_t& ent = e->h_block_table[(size_t)slot * e->max_blocks + blk];
        const bool need = blk * ps < len;
        if (need && ent < 0) {
            if (e->free_pages.empty()) { set_error("KV pool exhausted"); return WF_ERR_INVALID; }
            ent = e->free_pages.back();
            e->free_pages.pop_back();
        } else if (!need && ent >= 0) {
            e->free_pages.push_back(ent);
            ent = -1;
        }
    }
    e->table_dirty = true;
    e->h_pos[slot] = len;
    // on the engine stream (it is cudaStreamNonBlocking: a default-stream copy could race an in-flight asynchronous step, whose
    // advance kernel updates the same words)
    const int32_t pl[2] = {len, len + 1};
    if (cuda_check(cudaMemcpyAsync(e->d_pos + slot, &pl[0], 4, cudaMemcpyHostToDevice, e->stream), "ctx pos")) return WF_ERR_CUDA;
    if (cuda_check(cudaMemcpyAsync(e->d_len + slot, &pl[1], 4, cudaMemcpyHostToDevice, e->stream), "ctx len")) return WF_ERR_CUDA;
    if (cuda_check(cudaStreamSynchronize(e->stream), "ctx sync")) return WF_ERR_CUDA;
    if (fill_random) {
        // a page is one contiguous run of layers*2*Hkv*page_size*D elements
        const size_t page_elems = (size_t)e->cfg.layers * 2 * e->Hkv * ps * e->D;
        for (int blk = 0; blk * ps < len; ++blk) {
            const int32_t page = e->h_block_table[(size_t)slot * e->max_blocks + blk];
            int rc;
            if (e->kv_scales) {   // random int8 codes (the bytes of random bf16 values) and scales of N(1, 0.1)/42 so that values are O(1)
                rc = launch_fill_bf16(reinterpret_cast<unsigned char*>(e->pool) + (size_t)page * page_elems, page_elems / 2, 0xC0FFEEull + (uint64_t)page, 0.0f, 1.0f, e->stream);
                if (!rc) rc = launch_fill_f32(e->kv_scales + (size_t)page * (page_elems / e->D), page_elems / e->D, 1.0f / 42.0f, e->stream);
            } else {
                rc = launch_fill_bf16(e->pool + (size_t)page * page_elems, page_elems, 0xC0FFEEull + (uint64_t)page, 0.0f, 1.0f, e->stream);
            }
            if (rc) return rc;
        }
        return cuda_check(cudaStreamSynchronize(e->stream), "fill pool");
    }
    return WF_OK;
}

extern "C" int wf_engine_decode_device(wf_engine* e, int B, int use_graph) {
    if (!e || B < 1 || B > e->cfg.max_batch) { set_error("wf_engine_decode_device: bad batch %d", B); return WF_ERR_INVALID; }
    if (!e->finalized) { set_error("engine weights not finalized"); return WF_ERR_INVALID; }
    if (!e->tp_attached) { set_error("tensor-parallel engine: wf_engine_tp_attach has not been called"); return WF_ERR_INVALID; }
    int rc = prepare_pages(e, B);
    if (rc) return rc;
    rc = step_graph(e, B, /*feed_back=*/1, use_graph);
    if (rc) return rc;
    for (int b = 0; b < B; ++b) e->h_pos[b] += 1;
    return WF_OK;
}

extern "C" int wf_engine_decode(wf_engine* e, const int32_t* host_tokens, int B, int32_t* host_next_tokens) {
    if (!e || !host_tokens || !host_next_tokens || B < 1 || B > e->cfg.max_batch) { set_error("wf_engine_decode: bad arguments"); return WF_ERR_INVALID; }
    if (!e->finalized) { set_error("engine weights not finalized"); return WF_ERR_INVALID; }
    if (!e->tp_attached) { set_error("tensor-parallel engine: wf_engine_tp_attach has not been called"); return WF_ERR_INVALID; }
    int rc = prepare_pages(e, B);
    if (rc) return rc;
    memcpy(e->h_tok_pinned, host_tokens, (size_t)B * 4);
    if (cuda_check(cudaMemcpyAsync(e->d_tok, e->h_tok_pinned, (size_t)B * 4, cudaMemcpyHostToDevice, e->stream), "tokens H2D")) return WF_ERR_CUDA;
    rc = step_graph(e, B, /*feed_back=*/0, 1);
    if (rc) return rc;
    if (cuda_check(cudaMemcpyAsync(e->h_next_pinned, e->d_next, (size_t)B * 4, cudaMemcpyDeviceToHost, e->stream), "next D2H")) return WF_ERR_CUDA;
    if ((rc = check_device_error(e))) return rc;
    memcpy(host_next_tokens, e->h_next_pinned, (size_t)B * 4);
    for (int b = 0; b < B; ++b) e->h_pos[b] += 1;
    return WF_OK;
}

// ---- tensor parallel attach: one process per GPU, comm buffers exchanged as CUDA IPC handles ------------------------
extern "C" size_t wf_engine_tp_handle_bytes(void) { return sizeof(cudaIpcMemHandle_t); }

extern "C" int wf_engine_tp_export(wf_engine* e, void* handle_out) {
    if (!e || !handle_out) { set_error("wf_engine_tp_export: null argument"); return WF_ERR_INVALID; }
    cudaIpcMemHandle_t h;
    if (cuda_check(cudaIpcGetMemHandle(&h, e->comm), "cudaIpcGetMemHandle")) return WF_ERR_CUDA;
    memcpy(handle_out, &h, sizeof h);
    return WF_OK;
}

extern "C" int wf_engine_tp_attach(wf_engine* e, const void* handles) {
    if (!e || !handles) { set_error("wf_engine_tp_attach: null argument"); return WF_ERR_INVALID; }
    const int T = e->tp.size;
    for (int r = 0; r < T; ++r) {
        if (r == e->tp.rank) continue;
        if (e->peer_opened[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)handles + (size_t)r * sizeof h, sizeof h);
        void* p = nullptr;
        if (cuda_check(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle")) return WF_ERR_CUDA;
        e->peer_opened[r] = p;
        e->tp.peer[r] = (unsigned char*)p;
    }
    e->tp_attached = true;
    invalidate_graphs(e);
    return WF_OK;
}

extern "C" void* wf_engine_buffer(wf_engine* e, const char* name) {
    if (!e || !name) return nullptr;
    const std::string n(name);
    if (n == "hidden") return e->hid;
    if (n == "normed") return e->normed;
    if (n == "logits") return e->logits;
    if (n == "qkv") return e->qkv;
    if (n == "attn") return e->attn;
    if (n == "tokens") return e->d_tok;
    if (n == "next") return e->d_next;
    if (n == "positions") return e->d_pos;
    if (n == "seq_lens") return e->d_len;
    if (n == "pool") return e->pool;
    if (n == "block_table") return e->d_block_table;
    if (n == "q8") return e->q8;
    if (n == "trace") return e->trace;
    if (n == "acc_qkv") return e->acc_qkv;
    if (n == "acc_o") return e->acc_o;
    if (n == "acc_gu") return e->acc_gu;
    if (n == "acc_down") return e->acc_down;
    return nullptr;
}

// Times one component of the step in isolation on the engine's resident state (CUDA events on the
// engine stream; `iters` launches cycling over the layers so the working set exceeds L2).
extern "C" int wf_engine_time_kernel(wf_engine* e, const char* which, int B, int iters, float* ms_out) {
    if (!e || !which || !ms_out || B < 1 || B > e->cfg.max_batch || iters < 1) { set_error("wf_engine_time_kernel: bad arguments"); return WF_ERR_INVALID; }
    if (e->tp.size > 1) { set_error("wf_engine_time_kernel: single-GPU engines only"); return WF_ERR_UNSUPPORTED; }
    int rc = prepare_pages(e, B);
    if (rc) return rc;
    e->attn_seq_hint = seq_hint_for(e, B);
    const std::string w(which);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const wf_model_config& c = e->cfg;
    // "attn" times the kernel the step really runs.  Fused engines: paged_decode_attn_kernel<FUSED> (dequant of q|k|v from the raw
    // accumulator + RoPE + KV append + attention + in-kernel chunk merge) on a stand-in accumulator of random integers;
    // "attn_unfused" is the per-op kernel + attn_merge_kernel on a zeroed q.
    const bool fused_attn = (w == "attn") && e->fused;
    if (fused_attn) {
        if ((rc = launch_fill_randint_f32(e->acc_qkv, (size_t)B * e->qkv_rows, 0xA77E, 3000, e->stream))) return rc;
        if ((rc = launch_fill_f32(e->lin_asc, (size_t)B, 0.01f, e->stream))) return rc;
        if (cuda_check(cudaMemsetAsync(e->lin_rs, 0, (size_t)B * 4, e->stream), "rs memset")) return WF_ERR_CUDA;
        if (cuda_check(cudaMemsetAsync(e->merge_cnt, 0, (size_t)B * e->Hkv * 4, e->stream), "cnt memset")) return WF_ERR_CUDA;
    } else if (w == "attn" || w == "attn_unfused") {
        if (cuda_check(cudaMemsetAsync(e->qrot, 0, (size_t)B * e->attn_dim * 2, e->stream), "qrot memset")) return WF_ERR_CUDA;
    }
    auto body = [&](int it) -> int {
        const int l = it % c.layers;
        LayerW& L = e->layers[l];
        if (fused_attn) {
            FusedAttnArgs a = {};
            a.pool = e->pool; a.kv_scales = e->kv_scales; a.layers = c.layers; a.layer = l; a.Hkv = e->Hkv; a.page_size = c.page_size;
            a.block_table = e->d_block_table; a.max_blocks = e->max_blocks; a.seq_lens = e->d_len; a.positions = e->d_pos;
            a.B = B; a.H = e->H; a.scale = 1.0f / sqrtf((float)e->D); a.out = e->attn;
            a.n_out = 1; a.out_ld = e->attn_full; a.out_col0 = 0; a.signal_flag = -1; a.tp = e->tp;
            a.qkv_acc = e->acc_qkv; a.qkv_ld = e->qkv_rows; a.qkv_rowsum = e->lin_rs; a.qkv_ascale = e->lin_asc; a.qkv_wscale = L.scales;
            a.cos_tab = e->cos_tab; a.sin_tab = e->sin_tab; a.max_pos = c.max_seq;
            a.merge_cnt = e->merge_cnt;
            a.max_seq_hint = e->attn_seq_hint; a.ws = e->ws_attn; a.ws_bytes = e->ws_attn_bytes;
            return attn_decode_fused(a, e->stream, false);
        }
        if (w == "attn" || w == "attn_unfused")
            return attn_decode(e->qrot, e->pool, c.layers, l, e->Hkv, c.page_size, e->d_block_table, e->max_blocks, e->d_len, B, e->H,
                               1.0f / sqrtf((float)e->D), e->attn, WF_BF16, e->attn_seq_hint, e->ws_attn, e->ws_attn_bytes, e->stream, false, e->kv_scales);
        if (w == "qkv") { const int seg[3] = {e->H * e->D, (e->H + e->Hkv) * e->D, e->qkv_rows}; return linear(e, L.qkv, e->qkv_rows, e->hidden, B, L.scales, 3, seg, 0, e->qkv, e->qkv_rows, nullptr); }
        if (w == "o") { const int seg[1] = {e->hidden}; return linear(e, L.o, e->hidden, e->attn_dim, B, L.scales + 3, 1, seg, 0, e->normed, e->hidden, nullptr); }
        if (w == "gateup") { const int seg[2] = {e->ffn, 2 * e->ffn}; return linear(e, L.gu, 2 * e->ffn, e->hidden, B, L.scales + 4, 2, seg, WF_EPI_RELU2_MUL, e->gu, e->ffn, nullptr); }
        if (w == "down") { const int seg[1] = {e->hidden}; return linear(e, L.down, e->hidden, e->ffn, B, L.scales + 6, 1, seg, 0, e->normed, e->hidden, nullptr); }
        if (w == "linears") {
            int r;
            const int s3[3] = {e->H * e->D, (e->H + e->Hkv) * e->D, e->qkv_rows};
            const int s1[1] = {e->hidden};
            const int s2[2] = {e->ffn, 2 * e->ffn};
            if ((r = linear(e, L.qkv, e->qkv_rows, e->hidden, B, L.scales, 3, s3, 0, e->qkv, e->qkv_rows, nullptr))) return r;
            if ((r = linear(e, L.o, e->hidden, e->attn_dim, B, L.scales + 3, 1, s1, 0, e->normed, e->hidden, nullptr))) return r;
            if ((r = linear(e, L.gu, 2 * e->ffn, e->hidden, B, L.scales + 4, 2, s2, WF_EPI_RELU2_MUL, e->gu, e->ffn, nullptr))) return r;
            return linear(e, L.down, e->hidden, e->ffn, B, L.scales + 6, 1, s1, 0, e->normed, e->hidden, nullptr);
        }
        if (w == "lm_head") return launch_lm_head_argmax(e->normed, e->embed, B, c.vocab, e->hidden, nullptr, e->d_next, e->ws_lm, e->stream);
        if (w == "norm") return launch_rmsnorm_quant(e->hid, L.g_in, c.eps, B, e->hidden, nullptr, e->q8, e->a_scale, e->rowsum, e->stream, false);
        set_error("wf_engine_time_kernel: unknown component '%s'", which);
        return WF_ERR_INVALID;
    };
    for (int it = 0; it < 3; ++it) if ((rc = body(it))) return rc;
    cudaStreamSynchronize(e->stream);
    cudaEventRecord(e0, e->stream);
    for (int it = 0; it < iters; ++it) if ((rc = body(it))) return rc;
    cudaEventRecord(e1, e->stream);
    if (cuda_check(cudaEventSynchronize(e1), "time_kernel sync")) return WF_ERR_CUDA;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (fused_attn) {   // the step expects a zero accumulator
        if (cuda_check(cudaMemsetAsync(e->acc_qkv, 0, (size_t)B * e->qkv_rows * 4, e->stream), "acc memset")) return WF_ERR_CUDA;
        if (cuda_check(cudaStreamSynchronize(e->stream), "sync")) return WF_ERR_CUDA;
    }
    *ms_out = ms / iters;
    return WF_OK;
}
