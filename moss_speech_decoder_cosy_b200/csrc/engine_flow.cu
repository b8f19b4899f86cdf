// Engine: weight store, flow layout, estimator forward and the CFG Euler solver.
#include "engine.h"

#include <math.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>

namespace mossdec {

#define MD_CUDA(x)                                                                          \
    do {                                                                                    \
        cudaError_t e_ = (x);                                                               \
        if (e_ != cudaSuccess) return fail(std::string(#x) + ": " + cudaGetErrorString(e_)); \
    } while (0)

bool DeviceBuffer::reserve(size_t bytes) {
    if (bytes <= bytes_) return false;
    release();   // cudaFree waits for all work in flight on the device
    if (cudaMalloc(&ptr_, bytes) != cudaSuccess) {
        ptr_ = nullptr;
        bytes_ = 0;
        return true;
    }
    // (the legacy-stream memset is not ordered against non-blocking caller streams: finish it before anyone launches)
    cudaMemset(ptr_, 0, bytes);
    cudaDeviceSynchronize();
    bytes_ = bytes;
    return true;
}
void DeviceBuffer::release() {
    if (ptr_) cudaFree(ptr_);
    ptr_ = nullptr;
    bytes_ = 0;
}

PinnedStager::~PinnedStager() {
    for (auto& sl : slots_) {
        if (sl.ev) cudaEventDestroy(sl.ev);
        if (sl.p) cudaFreeHost(sl.p);
    }
}
void* PinnedStager::acquire(size_t bytes) {
    cur_ = (cur_ + 1) % kSlots;
    Slot& sl = slots_[cur_];
    if (sl.pending) {   // the copy that last read this slot (kSlots calls ago) must be done
        cudaEventSynchronize(sl.ev);
        sl.pending = false;
    }
    if (sl.cap < bytes) {
        if (sl.p) cudaFreeHost(sl.p);
        sl.p = nullptr;
        sl.cap = 0;
        const size_t cap = (bytes + 65535) & ~size_t(65535);
        if (cudaMallocHost(&sl.p, cap) != cudaSuccess) {
            sl.p = nullptr;
            return nullptr;
        }
        sl.cap = cap;
    }
    if (!sl.ev && cudaEventCreateWithFlags(&sl.ev, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    return sl.p;
}
void PinnedStager::commit(cudaStream_t st) {
    if (cur_ < 0) return;
    Slot& sl = slots_[cur_];
    if (cudaEventRecord(sl.ev, st) == cudaSuccess) sl.pending = true;
    else cudaStreamSynchronize(st);
}

Engine::Engine(const EngineConfig& cfg) : cfg_(cfg) {
    cudaGetDevice(&device_);
    cudaDeviceGetAttribute(&num_sms_, cudaDevAttrMultiProcessorCount, device_);
    // per-device kernel setup (opt-in shared memory sizes, constant tables) for THIS engine's device
    gemm_tc_init_device();
    ffn_tc_init_device();
    attention_ws_init_device();
    hift_init_device();
    const char* dv = getenv("MOSSDEC_DUAL");
    dual_stream_ = !(dv != nullptr && dv[0] == '0');
    if (dual_stream_) {
        if (cudaStreamCreateWithFlags(&st2_, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&ev_fork_, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&ev_join_, cudaEventDisableTiming) != cudaSuccess)
            dual_stream_ = false;
    }
    const char* fv = getenv("MOSSDEC_FUSE_OPROJ");
    fuse_oproj_ = !(fv != nullptr && fv[0] == '0');
    const char* gv = getenv("MOSSDEC_GRAPH");
    if (gv != nullptr && gv[0] >= '0' && gv[0] <= '2') graph_mode_ = gv[0] - '0';
    if (graph_mode_ && cudaStreamCreateWithFlags(&cap_st_, cudaStreamNonBlocking) != cudaSuccess) graph_mode_ = 0;
}

Engine::~Engine() {
    DeviceGuard g(device_);
    cudaDeviceSynchronize();
    drop_graphs();
    if (st2_) cudaStreamDestroy(st2_);
    if (cap_st_) cudaStreamDestroy(cap_st_);
    if (ev_fork_) cudaEventDestroy(ev_fork_);
    if (ev_join_) cudaEventDestroy(ev_join_);
    if (h_mask_lens_) cudaFreeHost(h_mask_lens_);
    if (d_mask_lens_) cudaFree(d_mask_lens_);
    for (auto& ev : prof_pool_) cudaEventDestroy(ev);
    for (auto& kv : tensors_)
        if (kv.second.ptr) cudaFree(kv.second.ptr);
}

int Engine::set_tensor(const char* name, const void* host, int dtype, int ndim, const int64_t* shape) {
    DeviceGuard g(device_);
    drop_graphs();            // captured launches hold weight pointers
    solve_tbias_steps_ = 0;
    enc_tmax_cached_ = -1;
    DevTensor t;
    t.dtype = dtype;
    size_t n = 1;
    for (int i = 0; i < ndim; ++i) {
        t.shape.push_back(shape[i]);
        n *= (size_t)shape[i];
    }
    t.bytes = n * (dtype == 1 ? 2 : 4);
    auto it = tensors_.find(name);
    if (it != tensors_.end()) {
        cudaFree(it->second.ptr);
        tensors_.erase(it);
    }
    // pad allocations so TMA boxes that start inside the tensor never touch unmapped memory
    MD_CUDA(cudaMalloc(&t.ptr, t.bytes + 1024));
    MD_CUDA(cudaMemset(t.ptr, 0, t.bytes + 1024));
    MD_CUDA(cudaMemcpy(t.ptr, host, t.bytes, cudaMemcpyHostToDevice));
    tensors_[name] = t;
    return 0;
}

const DevTensor* Engine::find(const std::string& name) {
    auto it = tensors_.find(name);
    if (it == tensors_.end()) {
        if (missing_.size() < 400) missing_ += name + " ";
        return nullptr;
    }
    return &it->second;
}

cudaEvent_t Engine::prof_event() {
    if (!prof_pool_.empty()) {
        cudaEvent_t e = prof_pool_.back();
        prof_pool_.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
void Engine::prof_begin(int cls, double flops, cudaStream_t st, uint64_t key) {
    ProfRec r{prof_event(), prof_event(), cls, flops, key};
    cudaEventRecord(r.a, st);
    prof_.push_back(r);
}
void Engine::prof_end(cudaStream_t st) { cudaEventRecord(prof_.back().b, st); }

int Engine::profile_read(double* out) {
    DeviceGuard g(device_);
    for (int i = 0; i < 9; ++i) out[i] = 0.0;
    if (cudaDeviceSynchronize() != cudaSuccess) return fail("profile_read: sync failed");
    for (auto& r : prof_) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.a, r.b);
        out[r.cls * 3 + 0] += ms;
        out[r.cls * 3 + 1] += 1.0;
        out[r.cls * 3 + 2] += r.flops;
        auto& acc = prof_by_key_[r.key];
        acc.first += ms;
        acc.second += 1;
        prof_pool_.push_back(r.a);
        prof_pool_.push_back(r.b);
    }
    prof_.clear();
    for (auto& r : bw_prof_) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.a, r.b);
        BwAcc& a = bw_acc_[r.name];
        a.ms += ms;
        a.n += 1;
        a.bytes += r.bytes;
        prof_pool_.push_back(r.a);
        prof_pool_.push_back(r.b);
    }
    bw_prof_.clear();
    if (getenv("MOSSDEC_PROFILE_DUMP")) {  // per GEMM shape/epilogue breakdown (developer aid)
        for (auto& kv : prof_by_key_) {
            const uint64_t k = kv.first;
            fprintf(stderr, "[mossdec prof] M=%-8llu N=%-5llu K=%-5llu act1=%llu ln1=%llu ln2=%llu resid=%llu f32=%llu bf16=%llu  n=%-6ld ms=%.3f\n",
                    (unsigned long long)(k >> 40), (unsigned long long)(k & 0xfff), (unsigned long long)((k >> 12) & 0xfff) * 64,
                    (unsigned long long)((k >> 24) & 7), (unsigned long long)((k >> 27) & 1), (unsigned long long)((k >> 28) & 1),
                    (unsigned long long)((k >> 29) & 1), (unsigned long long)((k >> 30) & 1), (unsigned long long)((k >> 31) & 1),
                    kv.second.second, kv.second.first);
        }
    }
    prof_by_key_.clear();
    return 0;
}

int Engine::profile_named(const char* name, double* out3) {
    auto it = bw_acc_.find(name);
    out3[0] = out3[1] = out3[2] = 0.0;
    if (it == bw_acc_.end()) return 0;
    out3[0] = it->second.ms;
    out3[1] = it->second.n;
    out3[2] = it->second.bytes;
    bw_acc_.erase(it);
    return 0;
}

int Engine::gemm(const GemmProblem& p, cudaStream_t st) {
    if (profile_) {
        int kb = 0;
        for (int s = 0; s < p.num_segs; ++s) kb += p.segs[s].nkb;
        const uint64_t key = (uint64_t)(p.N & 0xfff) | ((uint64_t)(kb & 0xfff) << 12) | ((uint64_t)(p.epi.act1 & 7) << 24) |
                             ((uint64_t)(p.epi.ln1_g != nullptr) << 27) | ((uint64_t)(p.epi.ln2_g != nullptr || p.epi.ln2_plain) << 28) |
                             ((uint64_t)(p.epi.resid != nullptr) << 29) | ((uint64_t)(p.epi.out_f32 != nullptr) << 30) |
                             ((uint64_t)(p.epi.out_bf16 != nullptr) << 31) | ((uint64_t)p.M << 40);
        prof_begin(0, 2.0 * p.M * p.N * kb * kGemmBK, st, key);
    }
    cudaError_t e = gemm_tc_launch(p, num_sms_, st);
    if (profile_) prof_end(st);
    ++launches_;
    if (e != cudaSuccess) return fail(std::string("gemm_tc_launch: ") + cudaGetErrorString(e));
    return 0;
}

int Engine::ffn(const FfnProblem& f, cudaStream_t st) {
    if (cfg_.channels != 256) return fail("fused FFN is built for channels == 256");
    if (profile_) prof_begin(2, 2.0 * f.M * 256.0 * (1024.0 * 2.0 + (f.att ? 512.0 : 0.0)), st, ((uint64_t)f.M << 40) | 0xfffu);
    cudaError_t e = ffn_tc_launch(f, num_sms_, st);
    if (profile_) prof_end(st);
    ++launches_;
    if (e != cudaSuccess) return fail(std::string("ffn_tc_launch: ") + cudaGetErrorString(e));
    return 0;
}

static std::string S(const char* fmt, int a = 0, int b = 0) {
    char buf[128];
    snprintf(buf, sizeof buf, fmt, a, b);
    return buf;
}

int Engine::finalize() {
    DeviceGuard g(device_);
    missing_.clear();
    const int ngroups = cfg_.n_mid + 2;
    const char* once[] = {"est.time.w1", "est.time.b1", "est.time.w2", "est.time.b2", "est.tbias.w", "est.tbias.b",
                          "est.down.conv.w", "est.down.conv.b", "est.up.conv.w", "est.up.conv.b", "est.final.w",
                          "est.final.b", "est.final.ln.g", "est.final.ln.b", "est.proj.w", "est.proj.b"};
    for (const char* n : once) find(n);
    const char* res[] = {"c1.w", "c1.b", "ln1.g", "ln1.b", "c2.w", "c2.b", "ln2.g", "ln2.b", "rc.w", "rc.b"};
    const char* tf[] = {"qkv.w", "qkv.b", "o.w", "o.b", "ff1.w", "ff1.b", "ff2.w", "ff2.b"};
    for (int i = 0; i < ngroups; ++i) {
        for (const char* n : res) find(S("est.res%d.", i) + n);
        for (int j = 0; j < cfg_.n_blocks; ++j)
            for (const char* n : tf) find(S("est.tf%d.%d.", i, j) + n);
    }
    if (!missing_.empty()) return fail("missing estimator tensors: " + missing_);
    finalized_ = true;
    finalize_hift();     // HiFT tensors are optional at load time; mossdec_hift fails loudly if they are absent
    finalize_encoder();  // likewise the encoder front end (mossdec_token2mel)
    return 0;
}

// ------------------------------------------------------------------------------------------------ layout
// Sequences are packed back to back with kGuard zero rows in front of each one; the CFG-unconditional half repeats
// the same layout rows_half further down (sequence ids nseq_half .. 2*nseq_half-1).
static constexpr int kGuard = 4;

int Engine::reserve_flow(size_t M, int nseq) {
    // layout arrays: [in_offset int64 x nseq | row_seq int32 x M | seq_start x nseq | seq_len x nseq]
    const size_t lay_bytes = (size_t)M * 4 + (size_t)nseq * (4 + 4 + 8) + 64;
    ws_layout_.reserve(lay_bytes + 256);
    if (!ws_layout_.as<void>()) return fail("cudaMalloc(layout) failed");
    // activation workspace (all buffers have M rows; one slack tile at the end for TMA boxes)
    const int C = cfg_.channels, H = cfg_.heads * cfg_.head_dim;
    const size_t Mr = M + kGemmBM;
    size_t need = 0;
    auto take = [&](size_t bytes) {
        need += (bytes + 1023) & ~size_t(1023);
    };
    take(Mr * cfg_.in_channels * 2); take(Mr * C * 2); take(Mr * C * 2); take(Mr * 3 * H * 2); take(Mr * H * 2);
    take(Mr * 4 * C * 2); take(Mr * 2 * C * 2); take(Mr * cfg_.n_feats * 4); take(Mr * C * 4); take(Mr * C * 4);
    take(Mr * cfg_.n_feats * 4);
    ws_flow_.reserve(need);
    if (!ws_flow_.as<void>()) return fail("cudaMalloc(flow workspace) failed");
    return 0;
}

int Engine::build_flow_layout(int nseq_half, const int32_t* lens, bool cfg_double, cudaStream_t st, int dense_T, int frame0) {
    int rows = 0, max_len = 0;
    const int nseq = cfg_double ? 2 * nseq_half : nseq_half;
    for (int s = 0; s < nseq_half; ++s) {
        if (lens[s] <= 0) return fail("sequence length must be positive");
        rows += kGuard + lens[s];
        max_len = lens[s] > max_len ? lens[s] : max_len;
    }
    rows += kGuard;
    const int rows_half = (rows + kGemmBM - 1) / kGemmBM * kGemmBM;
    const int M = cfg_double ? 2 * rows_half : rows_half;
    if (reserve_flow((size_t)M, nseq)) return -1;

    // build the layout arrays in a pinned slot (same order as on the device) and upload them with one async copy
    const size_t off_rowseq = ((size_t)nseq * 8 + 15) & ~size_t(15);
    const size_t lay_bytes = off_rowseq + (size_t)M * 4 + (size_t)nseq * 8;
    uint8_t* hp = static_cast<uint8_t*>(stager_.acquire(lay_bytes));
    if (!hp) return fail("cudaMallocHost(layout staging) failed");
    int64_t* h_off = reinterpret_cast<int64_t*>(hp);
    int32_t* h_rowseq = reinterpret_cast<int32_t*>(hp + off_rowseq);
    int32_t* h_start = h_rowseq + M;
    int32_t* h_len = h_start + nseq;
    for (int i = 0; i < M; ++i) h_rowseq[i] = -1;
    int64_t off = 0;
    int r = 0;
    for (int s = 0; s < nseq_half; ++s) {
        r += kGuard;
        h_start[s] = r;
        h_len[s] = lens[s];
        h_off[s] = dense_T > 0 ? (int64_t)s * cfg_.n_feats * dense_T : off;
        off += (int64_t)(lens[s] + frame0) * cfg_.n_feats;   // the ragged block also holds the frame0 frames that have no rows
        for (int t = 0; t < lens[s]; ++t) h_rowseq[r + t] = s;
        r += lens[s];
    }
    if (cfg_double) {
        for (int s = 0; s < nseq_half; ++s) {
            h_start[nseq_half + s] = h_start[s] + rows_half;
            h_len[nseq_half + s] = lens[s];
            h_off[nseq_half + s] = h_off[s];
            for (int t = 0; t < lens[s]; ++t) h_rowseq[rows_half + h_start[s] + t] = nseq_half + s;
        }
    }
    flow_M_ = M;
    flow_rows_half_ = rows_half;
    flow_nseq_ = nseq;
    flow_max_len_ = max_len;

    uint8_t* lp = ws_layout_.as<uint8_t>();
    MD_CUDA(cudaMemcpyAsync(lp, hp, lay_bytes, cudaMemcpyHostToDevice, st));
    stager_.commit(st);
    int32_t* d_rowseq = reinterpret_cast<int32_t*>(lp + off_rowseq);
    layout_.n_feats = cfg_.n_feats;
    layout_.rows_half = rows_half;
    layout_.noise_cstride = cfg_.noise_len;
    layout_.time_major = 0;
    layout_.dense_T = dense_T;
    layout_.frame0 = frame0;
    layout_.row_seq = d_rowseq;
    layout_.seq_start = d_rowseq + M;
    layout_.seq_len = d_rowseq + M + nseq;
    layout_.in_offset = reinterpret_cast<int64_t*>(lp);

    // activation workspace views (same carving as reserve_flow)
    const int C = cfg_.channels, H = cfg_.heads * cfg_.head_dim;
    const size_t Mr = (size_t)M + kGemmBM;
    size_t need = 0;
    auto take = [&](size_t bytes) {
        size_t o = need;
        need += (bytes + 1023) & ~size_t(1023);
        return o;
    };
    const size_t o_xin = take(Mr * cfg_.in_channels * 2), o_xn = take(Mr * C * 2), o_h1 = take(Mr * C * 2),
                 o_qkv = take(Mr * 3 * H * 2), o_att = take(Mr * H * 2), o_ff = take(Mr * 4 * C * 2),
                 o_cat = take(Mr * 2 * C * 2), o_xs = take(Mr * cfg_.n_feats * 4), o_hres = take(Mr * C * 4),
                 o_rres = take(Mr * C * 4), o_d = take(Mr * cfg_.n_feats * 4);
    uint8_t* b = ws_flow_.as<uint8_t>();
    xin_ = (__nv_bfloat16*)(b + o_xin);
    xn_ = (__nv_bfloat16*)(b + o_xn);
    h1_ = (__nv_bfloat16*)(b + o_h1);
    qkv_ = (__nv_bfloat16*)(b + o_qkv);
    att_ = (__nv_bfloat16*)(b + o_att);
    ff_ = (__nv_bfloat16*)(b + o_ff);
    cat_ = (__nv_bfloat16*)(b + o_cat);
    xstate_ = (float*)(b + o_xs);
    hres_ = (float*)(b + o_hres);
    rres_ = (float*)(b + o_rres);
    dout_ = (float*)(b + o_d);
    return 0;
}

int Engine::time_bias(const float* t_dev, int nt, float* temb, float* tbias, cudaStream_t st) {
    const int ngroups = cfg_.n_mid + 2;
    launch_time_embed(t_dev, nt, cfg_.in_channels, cfg_.time_dim, W<float>("est.time.w1"), W<float>("est.time.b1"),
                      W<float>("est.time.w2"), W<float>("est.time.b2"), temb, st);
    launch_time_bias(temb, nt, cfg_.time_dim, W<float>("est.tbias.w"), W<float>("est.tbias.b"),
                     ngroups * cfg_.channels, tbias, st);
    launches_ += 2;
    return 0;
}

// ------------------------------------------------------------------------------------------------ estimator
// Input: xin_ [M, 320] bf16 (guard rows zero).  Output: dout_ [M, 80] fp32 (guard rows zero).
// Frozen-prefix cache hooks (pc != nullptr; engine_solve.cu explains the scheme).  Causal convs look two rows back:
//   record (mode 1): the rows of frames prefix - 2, prefix - 1 of every conv input are saved;
//   reuse  (mode 2): the sequences hold live frames only and the two guard rows in front of each are filled from the cache
//                    right before the conv reads them (every producing epilogue zeroes guard rows again).
int Engine::prefix_halo(const PrefixCtx* pc, int conv_id, __nv_bfloat16* buf, int64_t ld, int width, int nseq, cudaStream_t st) {
    if (pc == nullptr || pc->mode == 0) return 0;
    __nv_bfloat16* cache = px_halo(pc->step, pc->branch, conv_id);
    if (pc->mode == 1) launch_prefix_halo(1, buf, ld, width, layout_.seq_start, pc->prefix - 2, px_d_slot_, nseq, cache, st);
    else launch_prefix_halo(0, buf, ld, width, layout_.seq_start, -2, px_d_slot_, nseq, cache, st);
    ++launches_;
    return 0;
}

int Engine::estimator_forward(int M, int nseq, int max_len, const float* tbias, int tbias_ld, int chunk,
                              cudaStream_t st, int row0, const PrefixCtx* pc) {
    const int C = cfg_.channels, H = cfg_.heads * cfg_.head_dim;
    const int ngroups = cfg_.n_mid + 2;
    const int nb = cfg_.n_blocks;
    const int32_t* row_seq = layout_.row_seq + row0;
    // workspace views of the rows [row0, row0 + M) (these locals shadow the members on purpose; row0 is a multiple of
    // 128, so the tiled fp32 matrices are offset by whole tiles).  Sequence starts stay relative to the view: the
    // second CFG half has the same layout as the first, so it uses the first half's seq_start / seq_len arrays.
    __nv_bfloat16* const xin_ = this->xin_ + (size_t)row0 * cfg_.in_channels;
    __nv_bfloat16* const xn_ = this->xn_ + (size_t)row0 * C;
    __nv_bfloat16* const h1_ = this->h1_ + (size_t)row0 * C;
    __nv_bfloat16* const qkv_ = this->qkv_ + (size_t)row0 * 3 * H;
    __nv_bfloat16* const att_ = this->att_ + (size_t)row0 * H;
    __nv_bfloat16* const cat_ = this->cat_ + (size_t)row0 * 2 * C;
    float* const hres_ = this->hres_ + (size_t)row0 * C;
    float* const rres_ = this->rres_ + (size_t)row0 * C;
    float* const dout_ = this->dout_ + (size_t)row0 * cfg_.n_feats;

    auto causal3 = [&](GemmProblem& p, const __nv_bfloat16* a, int cin) {
        p.a[0].ptr = a;
        p.a[0].rows = M;
        p.a[0].cols = cin;
        p.a[0].ld = cin;
        p.num_segs = 3;
        for (int k = 0; k < 3; ++k) p.segs[k] = {0, k - 2, 0, (cin + kGemmBK - 1) / kGemmBK};
        p.M = M;
    };
    auto dense = [&](GemmProblem& p, const __nv_bfloat16* a, int cin) {
        p.a[0].ptr = a;
        p.a[0].rows = M;
        p.a[0].cols = cin;
        p.a[0].ld = cin;
        p.num_segs = 1;
        p.segs[0] = {0, 0, 0, (cin + kGemmBK - 1) / kGemmBK};
        p.M = M;
    };

    const __nv_bfloat16* gin = xin_;  // input of the current group's resnet
    int gin_c = cfg_.in_channels;
    for (int g = 0; g < ngroups; ++g) {
        const std::string r = S("est.res%d.", g);
        if (g == ngroups - 1) {  // up block: cat(x, skip)
            gin = cat_;
            gin_c = 2 * C;
        }
        if (prefix_halo(pc, 2 * g, const_cast<__nv_bfloat16*>(gin), gin_c, gin_c, nseq, st)) return -1;
        {   // block1: causal conv -> LN -> Mish -> mask -> + time bias            (decoder.py:65-78, matcha decoder.py:56-57)
            GemmProblem p;
            causal3(p, gin, gin_c);
            p.w = W<__nv_bfloat16>(r + "c1.w");
            p.N = C; p.BN = C;
            p.epi.bias = W<float>(r + "c1.b");
            p.epi.ln1_g = W<float>(r + "ln1.g"); p.epi.ln1_b = W<float>(r + "ln1.b");
            p.epi.act1 = ACT_MISH;
            p.epi.row_seq = row_seq;
            p.epi.tbias = tbias + (size_t)g * C; p.epi.tbias_ld = tbias_ld;
            p.epi.out_bf16 = h1_; p.epi.ldo = C;
            if (gemm(p, st)) return -1;
        }
        {   // res_conv 1x1 on x*mask                                                 (matcha decoder.py:60)
            GemmProblem p;
            dense(p, gin, gin_c);
            p.w = W<__nv_bfloat16>(r + "rc.w");
            p.N = C; p.BN = C;
            p.epi.bias = W<float>(r + "rc.b");
            p.epi.out_f32 = rres_; p.epi.ldf = C; p.epi.f32_tiled = 1;
            if (gemm(p, st)) return -1;
        }
        if (prefix_halo(pc, 2 * g + 1, h1_, C, C, nseq, st)) return -1;
        {   // block2 + residual; also emits norm1 of the first transformer block
            GemmProblem p;
            causal3(p, h1_, C);
            p.w = W<__nv_bfloat16>(r + "c2.w");
            p.N = C; p.BN = C;
            p.epi.bias = W<float>(r + "c2.b");
            p.epi.ln1_g = W<float>(r + "ln2.g"); p.epi.ln1_b = W<float>(r + "ln2.b");
            p.epi.act1 = ACT_MISH;
            p.epi.row_seq = row_seq;
            p.epi.resid = rres_; p.epi.ldr = C;
            p.epi.out_f32 = hres_; p.epi.ldf = C; p.epi.f32_tiled = 1;
            p.epi.ln2_plain = 1;  // norm1 of the first transformer block (affine folded into its qkv weights)
            p.epi.out_bf16 = xn_; p.epi.ldo = C;
            if (gemm(p, st)) return -1;
        }
        for (int j = 0; j < cfg_.n_blocks; ++j) {
            const std::string t = S("est.tf%d.%d.", g, j);
            {   // q | k | v projection (q rows pre-scaled by 1/sqrt(d))
                GemmProblem p;
                dense(p, xn_, C);
                p.w = W<__nv_bfloat16>(t + "qkv.w");
                p.N = 3 * H; p.BN = 256;
                p.epi.bias = W<float>(t + "qkv.b");  // W . beta of the folded norm1
                p.epi.out_bf16 = qkv_; p.epi.ldo = 3 * H;
                if (gemm(p, st)) return -1;
            }
            if (profile_) prof_begin(1, 0.0, st);
            if (pc == nullptr || pc->mode == 0) {
                cudaError_t ae = launch_attention_ws(qkv_, M, att_, layout_.seq_start, layout_.seq_len, nseq, max_len,
                                                     cfg_.heads, chunk, st);
                if (ae != cudaSuccess) return fail(std::string("launch_attention_ws: ") + cudaGetErrorString(ae));
            } else {
                __nv_bfloat16* slab = px_kv(pc->step, pc->branch, g * nb + j);
                const int64_t slab_rows = (int64_t)px_slots_ * px_frames_ + 64;
                cudaError_t ae = cudaSuccess;
                if (pc->mode == 1) {
                    // record: k | v of the frozen frames -> cache; their queries see only frozen keys (chunk divides the
                    // prefix); the live queries run exactly as in reuse mode (frozen keys from the cache, same key tiling)
                    launch_prefix_kv_save(qkv_, H, layout_.seq_start, px_d_slot_, nseq, pc->prefix, slab, st);
                    ++launches_;
                    ae = launch_attention_ws(qkv_, M, att_, layout_.seq_start, px_d_len_pref_, nseq, pc->prefix, cfg_.heads, chunk, st);
                    ++launches_;
                    if (ae == cudaSuccess)
                        ae = launch_attention_ws_prefix(qkv_, M, att_, px_d_start_live_, px_d_len_live_, nseq, max_len - pc->prefix,
                                                        cfg_.heads, chunk, slab, slab_rows, pc->prefix, px_d_prow_, st);
                } else {
                    ae = launch_attention_ws_prefix(qkv_, M, att_, layout_.seq_start, layout_.seq_len, nseq, max_len, cfg_.heads,
                                                    chunk, slab, slab_rows, pc->prefix, px_d_prow_, st);
                }
                if (ae != cudaSuccess) return fail(std::string("launch_attention_ws (prefix): ") + cudaGetErrorString(ae));
            }
            if (profile_) prof_end(st);
            ++launches_;
            if (!fuse_oproj_) {   // out projection + residual, emits norm3 (separate launch: MOSSDEC_FUSE_OPROJ=0)
                GemmProblem p;
                dense(p, att_, H);
                p.w = W<__nv_bfloat16>(t + "o.w");
                p.N = C; p.BN = C;
                p.epi.bias = W<float>(t + "o.b");
                p.epi.row_seq = row_seq;
                p.epi.resid = hres_; p.epi.ldr = C;
                p.epi.out_f32 = hres_; p.epi.ldf = C; p.epi.f32_tiled = 1;
                p.epi.ln2_plain = 1;  // norm3 (affine folded into ff1)
                p.epi.out_bf16 = xn_; p.epi.ldo = C;
                if (gemm(p, st)) return -1;
            }
            {   // FF1 + erf-GELU + FF2 + residual in one kernel (hidden stays on chip); emits either the next block's
                // norm1 or the masked bf16 copy for the next conv
                FfnProblem f;
                f.x = xn_; f.M = M;
                if (fuse_oproj_) {   // attention out-projection + residual + norm3 as the prologue GEMM of the same kernel
                    f.att = att_;
                    f.wo = W<__nv_bfloat16>(t + "o.w");
                    f.bo = W<float>(t + "o.b");
                }
                f.w1 = W<__nv_bfloat16>(t + "ff1.w"); f.b1 = W<float>(t + "ff1.b");
                f.w2 = W<__nv_bfloat16>(t + "ff2.w"); f.b2 = W<float>(t + "ff2.b");
                f.hres = hres_;
                f.row_seq = row_seq;
                if (j + 1 < cfg_.n_blocks) {
                    f.ln = 1;  // norm1 of the next block (affine folded into its qkv weights)
                    f.out_bf16 = xn_; f.ldo = C;
                } else if (g == 0) {           // down block: skip connection -> right half of the concat buffer
                    f.out_bf16 = cat_ + C; f.ldo = 2 * C;
                } else if (g == ngroups - 2) { // last mid block -> left half of the concat buffer
                    f.out_bf16 = cat_; f.ldo = 2 * C;
                } else {                        // mid block -> next resnet input; up block -> upsample conv input
                    f.out_bf16 = xn_; f.ldo = C;
                }
                f.store_f32 = j + 1 < cfg_.n_blocks;  // the resnet that follows reads only the bf16 copy
                if (ffn(f, st)) return -1;
            }
        }
        if (g == 0) {  // "downsample" slot = plain causal conv when channels == [256]   (decoder.py:348-350)
            if (prefix_halo(pc, 2 * ngroups, cat_ + C, 2 * C, C, nseq, st)) return -1;
            GemmProblem p;
            p.a[0].ptr = cat_ + C; p.a[0].rows = M; p.a[0].cols = C; p.a[0].ld = 2 * C;
            p.num_segs = 3;
            for (int k = 0; k < 3; ++k) p.segs[k] = {0, k - 2, 0, C / kGemmBK};
            p.M = M;
            p.w = W<__nv_bfloat16>("est.down.conv.w");
            p.N = C; p.BN = C;
            p.epi.bias = W<float>("est.down.conv.b");
            p.epi.row_seq = row_seq;
            p.epi.out_bf16 = xn_; p.epi.ldo = C;  // first mid resnet input
            if (gemm(p, st)) return -1;
            gin = xn_;
            gin_c = C;
        } else if (g < ngroups - 1) {
            gin = xn_;
            gin_c = C;
        }
    }
    if (prefix_halo(pc, 2 * ngroups + 1, xn_, C, C, nseq, st)) return -1;
    {   // "upsample" slot = causal conv                                                (decoder.py:395-399)
        GemmProblem p;
        causal3(p, xn_, C);
        p.w = W<__nv_bfloat16>("est.up.conv.w");
        p.N = C; p.BN = C;
        p.epi.bias = W<float>("est.up.conv.b");
        p.epi.row_seq = row_seq;
        p.epi.out_bf16 = h1_; p.epi.ldo = C;
        if (gemm(p, st)) return -1;
    }
    if (prefix_halo(pc, 2 * ngroups + 2, h1_, C, C, nseq, st)) return -1;
    {   // final_block: conv -> LN -> Mish -> mask                                       (decoder.py:401, 492)
        GemmProblem p;
        causal3(p, h1_, C);
        p.w = W<__nv_bfloat16>("est.final.w");
        p.N = C; p.BN = C;
        p.epi.bias = W<float>("est.final.b");
        p.epi.ln1_g = W<float>("est.final.ln.g"); p.epi.ln1_b = W<float>("est.final.ln.b");
        p.epi.act1 = ACT_MISH;
        p.epi.row_seq = row_seq;
        p.epi.out_bf16 = xn_; p.epi.ldo = C;
        if (gemm(p, st)) return -1;
    }
    {   // final_proj 1x1 -> 80, masked                                                  (decoder.py:493-494)
        GemmProblem p;
        dense(p, xn_, C);
        p.w = W<__nv_bfloat16>("est.proj.w");
        p.N = cfg_.n_feats; p.BN = cfg_.n_feats;
        p.epi.bias = W<float>("est.proj.b");
        p.epi.row_seq = row_seq;
        p.epi.out_f32 = dout_; p.epi.ldf = cfg_.n_feats;
        if (gemm(p, st)) return -1;
    }
    return 0;
}

}  // namespace mossdec
