// Host-side context: the device image of one NeuralNetworkModel (parameters, optimizer state,
// per-step device state, workspace).  Opaque to callers of include/pozmlp.h.
#pragma once
#include "../../include/pozmlp.h"
#include "poz_common.cuh"
#include <string>
#include <vector>

struct poz_ctx {
    int device = 0;
    int nl = 0;                       // linear layers
    int has_bias = 0, loss = 0, dtype = 0;
    size_t esz = 4;                   // bytes per element of the master parameters
    std::vector<int> sizes, acts;
    std::vector<unsigned char> drop_z, drop_a;
    std::vector<long long> off;       // flat offsets of each layer's [W; b] image
    long long p_total = 0;

    size_t act_esz = 4;               // bytes per element of activations (2 in bf16 contexts)
    void* Pb = nullptr;               // bf16 copy of the parameters (bf16 contexts)
    float* colsum_scratch = nullptr;  // partial column sums for the bias gradient (bf16 contexts)
    void* wide_xg = nullptr;          // bf16 contexts: the sampled rows of a gathered batch, dense (inputs / targets)
    void* wide_tg = nullptr;
    long long wide_g_batch = 0;
    // split-bf16 tensor-core path of float32 contexts (csrc/poz_tc3.*)
    void* Wp = nullptr;               // weight planes, layer l at wp_off[l]: [3][in_l][wp_ld[l]] bf16
    std::vector<long long> wp_off, wp_ld;
    long long wp_total = 0;
    bool wp_stale = true;             // parameters changed since (some of) the planes were built
    std::vector<char> wp_fresh;       // per layer: planes match the current parameters
    // fp16-split inference (two planes, three MMAs per product; poz_tc3_gemm.cuh): its own weight planes
    void* Wp16 = nullptr;             // layer l at 2/3 of wp_off[l]: [2][in_l][wp_ld[l]] fp16
    bool w16_stale = true;            // parameters changed since these planes were built
    unsigned* w16_amax = nullptr;     // device: bits of max |W| per layer
    int* f16_flag = nullptr;          // device: set by a pass whose operands left the fp16-split range (the pass is repeated)
    bool tc3_f16 = true;              // POZ_TC3_F16=0 keeps large-batch inference on the split-bf16 kernels
    long long tc3_f16_min_rows = 16384;
    void* tc_tr_x = nullptr;          // training: planes of the (gathered) input batch
    void* tc_tr_dz = nullptr;         // training: planes of dZ of the first layer
    long long tc_tr_batch = 0;
    // measured at batch 1024 (C2) the plane splits cost what the tensor cores save, so the SIMT
    // kernels keep the first layer below this batch (POZ_TC3_TRAIN_MIN_BATCH overrides)
    long long tc3_train_min_batch = 2304;   // (measured, us per configs[1] step, cluster split-K path / this path: 2048 rows 95.7 / 100.6,
                                            //  2560 rows 153.2 / 119.5, 4096 rows 208.0 / 150.1)
    bool tc3_train = true;            // first layer of float32 training on tcgen05 when it is deep enough
    // small-batch tensor-core path of the first layer (csrc/poz_tcs_kernel.cuh): forward and weight gradient of
    // layer 0 as cluster split-K tcgen05 launches, the optimizer applied in the weight-gradient epilogue
    bool tcs_enabled = true;          // POZ_TCS=0 keeps the SIMT kernels
    bool tcs_fuse_update = true;      // POZ_TCS_FUSE_UPDATE=0: the weight-gradient launch stores the gradient instead
    long long tcs_min_batch = 96, tcs_max_batch = 2560;
    bool tcs_live = false;            // layer 0's weight planes are kept equal to its parameters at all times
    std::vector<void*> tcs_dzp;       // planes of dZ of layer l [3][batch][round8(out_l)] (layer 0 only unless grouped)
    long long tcs_dz_batch = 0;
    // All layers' weight gradients in one grouped launch: measured slower on the C2 stack (36 clusters of 4 CTAs where
    // 33 fit at once: a second wave), so opt-in: POZ_TCS_GROUP=1
    bool tcs_group = false;
    unsigned upd_done_mask = 0;       // layers whose update this step's backward already applied
    bool tail_fused = true;           // last hidden unit + head + their backward-x in one launch (POZ_TAIL=0 disables)
    long long tail_max_batch = 4096;
    bool pdl = true;                  // the launches of the narrow step's critical path as programmatic dependents (POZ_PDL=0: plain edges)
    int tail_cluster = 4;             // blocks of the tail launch that share the reads of W1 (POZ_TAIL_CLUSTER = 1, 2, 4)
    bool tail_group = true;           // + the weight gradients of layers 0 and 1 in one tensor-core launch, head / b1 updated from the
                                      //   tail's partial sums (POZ_TAIL_GROUP=0 disables)
    float* tail_part = nullptr;       // [blocks][stride] partial sums written by tail_fused_kernel
    long long tail_part_elems = 0;
    bool in_step = false;             // inside poz_train_step (backward + apply as one call: the update may be fused)
    void* tc_scr[3] = {nullptr, nullptr, nullptr};   // plane stacks: input chunk, ping, pong
    float* tc_logits = nullptr;       // fp32 logits of a chunk (trailing softmax)
    long long tc_chunk = 0;
    long long tc3_min_batch = 2048;   // smaller batches stay on the SIMT kernels (POZ_TC3_MIN_BATCH overrides)
    // general stacks: embedding head, flatten row multipliers, batchnorm (csrc/poz_general.inc)
    bool general = false;             // any of the above present -> the general path runs the stack
    int emb_vocab = 0, emb_dim = 0, emb_drop = 0, tokens = 1;
    long long off_emb = -1;           // flat offset of the embedding table
    std::vector<int> kin;             // input width of layer l (= sizes[l] unless a flatten sits in front of it)
    std::vector<int> rmul;            // rows of layer l = batch * rmul[l]
    std::vector<int> has_bn, kz, kbn, ka;
    std::vector<double> bn_eps, bn_mom;
    std::vector<long long> off_bn;    // flat offset of [gain; bias] of layer l, -1 if none
    std::vector<void*> bn_run;        // running [mean; variance] of layer l
    std::vector<void*> bn_stat;       // [mean; 1/std] used by the current step
    std::vector<void*> bn_sums;       // backward column sums
    double* bn_part = nullptr;        // partial column sums [BN_ROW_SPLITS][2][widest batchnorm]
    std::vector<void*> ybuf, dybuf, dy_rec;   // batchnorm output (diagnostics), its pre-dropout gradient
    void* ebuf = nullptr;             // embedding output [batch*tokens][emb_dim]
    void* debuf = nullptr;            // its gradient
    void* de_rec = nullptr;           // diagnostics: gradient of the recorded embedding output
    long long ws_gen = 0;
    void* P = nullptr;                // flat parameters
    void* AM = nullptr;               // Adam exp_avg
    void* AV = nullptr;               // Adam exp_avg_sq
    void* G = nullptr;                // gradient slabs [g_slabs][p_total]
    void* Gdiag = nullptr;            // finished gradients of the last diagnostics step
    int g_slabs = 0, max_splits = 16;
    std::vector<int> seg_splits;
    std::vector<cudaEvent_t> grad_ready;   // recorded when layer l's weight gradient is complete
    std::vector<cudaEvent_t> dz_ready;     // recorded (compute stream) when layer l's dZ is complete
    cudaStream_t side_stream = nullptr;    // weight-gradient GEMMs of layers >= 1 run here, beside the backward-x chain
    cudaStream_t side_stream2 = nullptr;   // ... alternating with this one, so that two of them do not queue behind each other
    bool fork_dw = true;
    cudaEvent_t sb_fork = nullptr, sb_done = nullptr;   // step bookkeeping forked beside the forward GEMMs

    poz::DevState* st = nullptr;
    double* cost_ring = nullptr;  long long cost_cap = 0;
    double* stats = nullptr;      long long stats_cap = 0;
    double* scalar_cost = nullptr;
    double* stage = nullptr;      size_t stage_elems = 0;

    // workspace
    long long ws_train = 0, ws_infer = 0, ws_diag = 0;
    std::vector<void*> zbuf, abuf, dzbuf, dz_rec, da_rec;
    void* ping = nullptr;
    void* pong = nullptr;

    struct ProfRec { std::string tag; cudaEvent_t a, b; };
    std::vector<ProfRec> prof;
    bool profiling = false;

    // CUDA-graph replay of whole train steps (same buffers, same batch): see poz_train_step
    struct StepGraph {
        const void* x; const void* t; const void* idx; long long batch;
        cudaGraphExec_t exec = nullptr;   // null until the key has been seen twice
        int launches = 0;
    };
    std::vector<StepGraph> graphs;
    std::vector<StepGraph> fwd_graphs;   // inference calls: x = input, t = output buffer
    bool graphs_enabled = true;
    int small_infer_cluster = 0;      // cluster size of the few-row inference kernel (0 = not probed, -1 = unavailable)
    bool fuse_head = true;            // narrow softmax head in one kernel (POZ_FUSE_HEAD=0 disables)

    double graph_sig[6] = {0, 0, 0, 0, 0, 0};

    // host-buffer stepping (poz_train_step_host): POZ_HOST_LAG + 1 device slots, copy + read-back streams
    struct HostFeed {
        static constexpr int SLOTS = POZ_HOST_LAG + 1;
        void* x[SLOTS] = {};
        void* t[SLOTS] = {};
        cudaEvent_t copied[SLOTS] = {}, copied_t[SLOTS] = {}, consumed[SLOTS] = {};
        long long epoch_of[SLOTS] = {};   // epoch index of the step that last used the slot
        cudaStream_t copy_stream = nullptr, t_stream = nullptr, read_stream = nullptr;   // inputs, targets, cost read-back
        double* cost_pinned = nullptr;
        long long batch = 0;           // capacity of the slots
        long long calls = 0;           // host steps since the last flush / poz_train_begin
    } feed;
    long long epoch_host = 0;          // steps launched since poz_train_begin (mirrors DevState::next_epoch)

    poz_train_config cfg;
    bool train_begun = false, training_pass = false;
    bool grads_pending = false, grads_reduced = false, diag_pending = false, have_diag = false;
    long long diag_batch = 0, last_batch = 0;
};
