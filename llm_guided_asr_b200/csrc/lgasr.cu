// Host side of the C ABI declared in include/lgasr.h (CUDA runtime only; no torch types).
#include "../../include/lgasr.h"

#include <atomic>
#include <climits>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "lgasr_kernels.cuh"
#include "lgasr_scorer_kernels.cuh"
#include "lgasr_step.cuh"
#include "lgasr_step2.cuh"
#include "lgasr_state_kernels.cuh"

using namespace lgasr;

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CK(call)                                                                                      \
  do {                                                                                                \
    cudaError_t e_ = (call);                                                                          \
    if (e_ != cudaSuccess) return fail(LGASR_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
  } while (0)

#define LAUNCH_CK()                                                                                   \
  do {                                                                                                \
    g_launches.fetch_add(1, std::memory_order_relaxed);                                               \
    cudaError_t e_ = cudaGetLastError();                                                              \
    if (e_ != cudaSuccess) return fail(LGASR_ECUDA, "%s:%d kernel launch: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
  } while (0)

struct lgasr_handle {
  lgasr_config cfg;
  Dev d;
  int pp, rp, step, begun, have_post;
  int refilled;  // lgasr_refill was used since lgasr_begin: the utterances are at different decode positions
  int NL;
  size_t smem_rows, smem_sel;
  StepLaunch sl;     // k_step launch geometry
  size_t smem_step;
  int ept;           // 8 or 16 elements per selection thread and chunk
  int fused;         // 1: one k_step launch per position; 0: k_score_rows + k_select
  int cluster;       // 1: k_step is launched as clusters of `beam` CTAs (records through distributed shared memory)
  int seq_ok;        // 1: the whole grid of k_step is one wave, replay positions use the SEQ instance
  int v2;            // 1: k_step2 is available for this handle (geometry computed)
  int v2_all;        // 1: every fused launch takes k_step2 (LGASR_V2=1); 0: only bf16 / normalised term matrices do
  StepLaunch sl2;    // k_step2 launch geometry
  size_t smem_step2;
  int cluster2;      // k_step2 launched as clusters of `beam` CTAs
  int nth2;          // threads per CTA of k_step2 (256 or 128)
  int mb2;           // its resident-CTA target per SM (launch bound): 5 (256 threads), 6, or 8 = the dense instance
  std::vector<void*> allocs;
  std::vector<long long> primer;
  std::vector<int> xlens;
  int* h_xlen_pin;  // pinned staging of lgasr_set_posteriors: xlens[B]
  cudaEvent_t ev_xlen;  // the copy out of h_xlen_pin has been performed
  int* d_status;    // device status word: bit 0 = the ended-hypotheses pool of some utterance overflowed
  int* h_pin;  // pinned staging: n_run[B], done[B]
  int* h_begin_pin;  // pinned staging of lgasr_begin: maxlen[B], minlen[B], primer (int64)
  cudaEvent_t ev_begin;  // the copies out of h_begin_pin have been performed
  long long* d_primer;
  float* d_logp_stage;  // staging for set_posteriors_host
  size_t logp_stage_elems;
  // host-side replay staging
  float* term_stage[2][LGASR_MAX_TERMS];
  size_t term_stage_elems;
  cudaStream_t copy_stream;
  cudaEvent_t ev_copy[2], ev_comp[2];
  // ended-pool cache
  bool cache_valid;
  int cache_max_n, cache_steps;
  std::vector<int> c_n_ended;
  std::vector<float> c_score, c_scores;
  std::vector<int4> c_meta;
  std::vector<int> c_bp_par, c_bp_tok;
  // per-kernel profiling
  int profiling;
  std::vector<cudaEvent_t> prof_ev;  // 3 per step
  // CUDA graph of the last device-resident replay (the launch-bound loop of lgasr_decode_replay)
  cudaGraphExec_t rg_exec;
  std::vector<const float*> rg_ptrs;
  std::vector<float> rg_cst, rg_w;
  long long rg_ld;
  int rg_steps, rg_pp, rg_rp, rg_step0;
  cudaStream_t rg_stream;  // the legacy default stream cannot be captured: graphs run on this one
  cudaEvent_t rg_ev[2];
};

template <typename T>
static int dalloc(lgasr_handle* h, T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, n * sizeof(T) + 16);
  if (e != cudaSuccess) return fail(LGASR_ENOMEM, "cudaMalloc(%zu bytes): %s", n * sizeof(T), cudaGetErrorString(e));
  h->allocs.push_back(q);
  *p = reinterpret_cast<T*>(q);
  return LGASR_OK;
}

#define TRY(x)            \
  do {                    \
    int rc_ = (x);        \
    if (rc_) return rc_;  \
  } while (0)

// instances of k_step2 for a CTA size: [0..3] single CTAs (general, one matrix, two matrices, one matrix + trace),
// [4..7] the same as thread-block clusters
template <int NTH, int MB>
static void step2_instances_of(const void** f) {
  f[0] = (const void*)k_step2<NTH, false, 0, false, MB>;
  f[1] = (const void*)k_step2<NTH, false, 1, false, MB>;
  f[2] = (const void*)k_step2<NTH, false, 2, false, MB>;
  f[3] = (const void*)k_step2<NTH, false, 1, true, MB>;
  f[4] = (const void*)k_step2<NTH, true, 0, false, MB>;
  f[5] = (const void*)k_step2<NTH, true, 1, false, MB>;
  f[6] = (const void*)k_step2<NTH, true, 2, false, MB>;
  f[7] = (const void*)k_step2<NTH, true, 1, true, MB>;
}
static void step2_instances(int nth, int mb, const void** f) {
  if (nth == 256) step2_instances_of<256, 5>(f);
  else if (mb == 8) step2_instances_of<128, 8>(f);
  else step2_instances_of<128, 6>(f);  // 64 threads per CTA were measured too (slower: profiles/README.md) and dropped
}

extern "C" int lgasr_version(void) { return LGASR_VERSION; }
extern "C" const char* lgasr_last_error(void) { return g_err; }
extern "C" int64_t lgasr_launch_count(void) { return g_launches.load(); }

extern "C" int lgasr_create(const lgasr_config* c, lgasr_handle** out) {
  if (!c || !out) return fail(LGASR_EINVAL, "null argument");
  if (c->n_utts <= 0 || c->max_frames <= 0 || c->vocab <= 1 || c->beam <= 0)
    return fail(LGASR_EINVAL, "bad extents B=%d T=%d V=%d beam=%d", c->n_utts, c->max_frames, c->vocab, c->beam);
  if (c->n_terms < 0 || c->n_terms > LGASR_MAX_TERMS) return fail(LGASR_EINVAL, "n_terms %d > %d", c->n_terms, LGASR_MAX_TERMS);
  if (c->n_terms == 0 && !c->use_ctc) return fail(LGASR_EINVAL, "no scorer at all");
  if (c->pre_beam_key >= c->n_terms) return fail(LGASR_EINVAL, "pre_beam_key %d out of range", c->pre_beam_key);
  if (c->eos < 0 || c->eos >= c->vocab || c->blank < 0 || c->blank >= c->vocab) return fail(LGASR_EINVAL, "bad eos/blank");
  if (c->primer_len < 1 || c->max_len_cap < 1) return fail(LGASR_EINVAL, "bad primer_len/max_len_cap");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(LGASR_ECUDA, "no CUDA device: this library has no CPU fallback");

  lgasr_handle* h = new (std::nothrow) lgasr_handle();
  if (!h) return fail(LGASR_ENOMEM, "host alloc");
  h->cfg = *c;
  Dev& d = h->d;
  memset(&d, 0, sizeof(d));
  d.B = c->n_utts;
  d.T = c->max_frames;
  d.Tp = (c->max_frames + 3) & ~3;
  d.V = c->vocab;
  d.W = c->beam;
  d.S = c->pre_beam;
  d.K = c->n_terms;
  d.key = c->pre_beam_key;
  d.use_ctc = c->use_ctc ? 1 : 0;
  d.w_ctc = c->ctc_weight;
  d.blank = c->blank;
  d.sos = c->sos;
  d.eos = c->eos;
  d.maxlen_cap = c->max_len_cap;
  d.Lcap = c->primer_len + c->max_len_cap + 1;
  d.Ecap = c->beam * c->max_len_cap;
  d.end_detect = c->end_detect ? 1 : 0;
  d.NS = d.K + d.use_ctc;
  d.P = c->primer_len;
  d.d_end = -10.0;  // float(np.log(1 * np.exp(-10))) == -10.0 exactly, e2e_asr_common.py:17
  // do_pre_beam rule of beam_search.py:115-119
  const bool do_pre_beam = d.use_ctc && d.key != -2 && d.S < d.V;
  d.mode = !d.use_ctc ? kModeNoCtc : (do_pre_beam ? kModePreBeam : kModeFullCtc);
  // the pre-beam key only exists while there is a pre-beam: without a partial scorer the reference ranks the full
  // weighted sum whatever pre_beam_score_key says (do_pre_beam False, beam_search.py:115-119; batch_beam_search.py:302-308)
  if (d.mode != kModePreBeam) d.key = -1;
  if (d.mode == kModePreBeam && d.S < 1) return fail(LGASR_EINVAL, "pre_beam must be >= 1");
  const int n_sel = d.mode == kModePreBeam ? d.S : d.W;
  if (n_sel > LGASR_MAX_PREBEAM) {
    delete h;
    return fail(LGASR_EINVAL, "selection width %d > %d unsupported", n_sel, LGASR_MAX_PREBEAM);
  }
  d.R = d.mode == kModePreBeam ? d.S + 1 : d.W;
  h->NL = (n_sel + 31) / 32;
  if (h->NL == 3) h->NL = 4;
  const int capS = 32 * h->NL;
  h->smem_rows = (size_t)4 * d.Tp * sizeof(float) + (size_t)kRowWarps * capS * 8 + (size_t)capS * 8;
  h->smem_sel = (size_t)utt_select_smem_floats(d.W, d.R, d.K + (c->use_ctc ? 1 : 0)) * sizeof(float);
  if (h->smem_rows > 220 * 1024 || h->smem_sel > 220 * 1024) {
    delete h;
    return fail(LGASR_EINVAL, "shared memory need too large (T=%d beam=%d)", d.T, d.W);
  }

  // ---- k_step geometry -------------------------------------------------------------------------
  {
    const char* env = getenv("LGASR_LEGACY");
    h->fused = (d.mode != kModeFullCtc) && !(env && env[0] == '1');
    StepLaunch& sl = h->sl;
    const int K = d.K > 0 ? d.K : 1;
    // stage budget: 6144 floats (24 KB).  A row set that fits is ONE chunk; longer rows are streamed through
    // two stages of half the budget each.
    int budget = 6144;
    const int V4 = (d.V + 3) & ~3;
    const char* ncl0 = getenv("LGASR_NO_CLUSTER");
    const bool want_cluster = h->fused && d.W >= 2 && d.W <= 16 && !(ncl0 && ncl0[0] == '1');
    const long long selw0 = utt_select_smem_floats(d.W, d.R, d.NS);
    // dynamic shared memory of a CTA for a given ring budget (the arena also holds the candidate rows and, in the
    // ticket instance, the selection's working set)
    auto total_for = [&](int bud) -> long long {
      long long ar = ((long long)K * V4 <= bud) ? (long long)K * V4 : 2LL * K * ((bud / (2 * K)) & ~3);
      long long cd = (long long)n_sel * d.Tp;
      if (cd > 6144) cd = 6144;
      if (cd < d.Tp) cd = d.Tp;
      if (d.mode == kModePreBeam && cd > ar) ar = cd;
      if (selw0 > ar) ar = selw0;
      ar = (ar + 3) & ~3LL;
      return (long long)sizeof(StepSmem) + 6LL * d.Tp * sizeof(float) + ar * (long long)sizeof(float) +
             (want_cluster ? selw0 * (long long)sizeof(float) : 0);
    };
    if ((long long)K * V4 > budget) {
      // a streamed row set: if the default ring is what pushes the CTA over the five-per-SM footprint AND that
      // turns a one-wave grid into two (two full scorers at configs[1] size: 45.4 KB -> four CTAs per SM -> the
      // 640 CTAs no longer fit, 37 us per launch instead of 24), shrink the ring by the few hundred bytes it
      // takes to stay at five.  Otherwise the larger chunks win (OWSM shape: 29 vs 36 us).
      const long long per_cta5 = (227LL * 1024) / 5 - 1024 - 512;  // 1 KB reserved per CTA + static shared memory
      int dev = 0, sms = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      const long long tot0 = total_for(budget);
      const long long per_sm0 = (227LL * 1024) / (tot0 + 1024 + 512);
      const long long ctas = (long long)d.B * d.W;
      if (tot0 > per_cta5 && ctas > per_sm0 * sms && ctas <= 5LL * sms) {
        const long long fixed = tot0 - ((tot0 - (long long)sizeof(StepSmem) - 6LL * d.Tp * sizeof(float) -
                                         (want_cluster ? selw0 * (long long)sizeof(float) : 0)));
        const long long room = (per_cta5 - fixed) / (long long)sizeof(float);
        if (room < budget && room >= 4096 && ((room / (2 * K)) & ~3LL) >= 256 && total_for((int)room) <= per_cta5)
          budget = (int)room;
      }
    }
    int nst = 1;
    if ((long long)K * V4 <= budget) {
      sl.CH = V4;
    } else {
      nst = 2;
      sl.CH = (budget / (2 * K)) & ~3;
    }
    sl.NST = nst;
    sl.nchunks = (d.V + sl.CH - 1) / sl.CH;
    sl.kk_full = (n_sel + kSelWarps - 1) / kSelWarps;
    sl.tma_rows = 0;  // decided per step from the pointers
    long long arena = (long long)nst * K * sl.CH;
    long long cand = (long long)n_sel * d.Tp;
    if (cand > 6144) cand = 6144;
    if (cand < d.Tp) cand = d.Tp;
    if (d.mode == kModePreBeam && cand > arena) arena = cand;
    const long long selw = utt_select_smem_floats(d.W, d.R, d.NS);
    if (selw > arena) arena = selw;
    arena = (arena + 3) & ~3LL;
    sl.arena_floats = (int)arena;
    int rpg = (int)(arena / d.Tp);
    if (rpg < 1) rpg = 1;
    if (rpg > n_sel) rpg = n_sel;
    sl.rows_per_group = rpg;
    const char* ncl = getenv("LGASR_NO_CLUSTER");
    h->cluster = h->fused && d.W >= 2 && d.W <= 16 && !(ncl && ncl[0] == '1');
    h->smem_step = sizeof(StepSmem) + (size_t)6 * d.Tp * sizeof(float) + (size_t)arena * sizeof(float);
    if (h->cluster) h->smem_step += (size_t)utt_select_smem_floats(d.W, d.R, d.NS) * sizeof(float);
    if (h->fused && h->smem_step > 220 * 1024) {
      delete h;
      return fail(LGASR_EINVAL, "shared memory need too large for the fused step (T=%d beam=%d K=%d)", d.T, d.W, d.K);
    }
  }

  const size_t BW = (size_t)d.B * d.W;
  int rc = 0;
  float* xT = nullptr;
  int* xlen = nullptr;
  if (d.use_ctc) {
    rc |= dalloc(h, &xT, (size_t)d.B * d.V * d.Tp);
    for (int p = 0; p < 2; ++p) rc |= dalloc(h, &d.r[p], BW * 2 * d.Tp);
  }
  rc |= dalloc(h, &xlen, d.B);
  d.xT = xT;
  d.xlen = xlen;
  for (int p = 0; p < 2; ++p) {
    rc |= dalloc(h, &d.s_prev[p], BW);
    rc |= dalloc(h, &d.score[p], BW);
    rc |= dalloc(h, &d.scores[p], BW * (d.NS > 0 ? d.NS : 1));
    rc |= dalloc(h, &d.yseq[p], BW * d.Lcap);
    rc |= dalloc(h, &d.run_slot[p], BW);
  }
  int *maxlen = nullptr, *minlen = nullptr;
  for (int p = 0; p < 2; ++p) {
    rc |= dalloc(h, &d.cur_len[p], d.B);
    rc |= dalloc(h, &d.n_run[p], d.B);
    rc |= dalloc(h, &d.done[p], d.B);
    rc |= dalloc(h, &d.rmeta[p], BW);
    rc |= dalloc(h, &d.urec[p], d.B);
    rc |= dalloc(h, &d.hrec[p], 2 * BW);
  }
  rc |= dalloc(h, &d.ustat, d.B);
  rc |= dalloc(h, &d.status, 1);
  rc |= dalloc(h, &d.arrive, d.B);
  rc |= dalloc(h, &d.step_flag, d.B);
  if (getenv("LGASR_TRACE")) {
    rc |= dalloc(h, &d.trace, BW * 32);
    if (!rc) cudaMemset(d.trace, 0, BW * 32 * sizeof(long long));
  }
  rc |= dalloc(h, &d.step_parent, BW);
  rc |= dalloc(h, &d.step_tok, BW);
  rc |= dalloc(h, &maxlen, d.B);
  rc |= dalloc(h, &minlen, d.B);
  d.maxlen = maxlen;
  d.minlen = minlen;
  rc |= dalloc(h, &d.cand_total, BW * d.R);
  rc |= dalloc(h, &d.cand_tok, BW * d.R);
  rc |= dalloc(h, &d.cand_psi, BW * d.R);
  rc |= dalloc(h, &d.bp_par, (size_t)d.B * d.maxlen_cap * d.W);
  rc |= dalloc(h, &d.bp_tok, (size_t)d.B * d.maxlen_cap * d.W);
  rc |= dalloc(h, &d.n_ended, d.B);
  rc |= dalloc(h, &d.end_score, (size_t)d.B * d.Ecap);
  rc |= dalloc(h, &d.end_meta, (size_t)d.B * d.Ecap);
  rc |= dalloc(h, &d.end_scores, (size_t)d.B * d.Ecap * (d.NS > 0 ? d.NS : 1));
  rc |= dalloc(h, &d.best_by_len, (size_t)d.B * (d.Lcap + 2));
  rc |= dalloc(h, &d.best_all, d.B);
  rc |= dalloc(h, &h->d_primer, c->primer_len);
  if (rc) {
    lgasr_destroy(h);
    return LGASR_ENOMEM;
  }
  if (cudaMallocHost(&h->h_pin, sizeof(int) * (2 * d.B + 2 * d.B * d.W + 1)) != cudaSuccess ||
      cudaMallocHost(&h->h_xlen_pin, sizeof(int) * d.B) != cudaSuccess ||
      cudaEventCreateWithFlags(&h->ev_xlen, cudaEventDisableTiming) != cudaSuccess) {
    lgasr_destroy(h);
    return fail(LGASR_ENOMEM, "cudaMallocHost");
  }
  h->xlens.assign(d.B, d.T);
  h->cache_valid = false;

  // opt in to large dynamic shared memory.  The attribute is per function AND per device, and another handle of the
  // same device may need more: it only ever grows, with one high-water mark per device.
  {
    static std::mutex mu;
    constexpr int kMaxDev = 64;
    static size_t hw_rows[kMaxDev][5] = {}, hw_sel[kMaxDev] = {}, hw_step[kMaxDev] = {};
    static bool hw_np[kMaxDev] = {};
    int dev_id = 0;
    cudaGetDevice(&dev_id);
    dev_id = dev_id < 0 || dev_id >= kMaxDev ? 0 : dev_id;
    size_t* cur_rows = hw_rows[dev_id];
    size_t& cur_sel = hw_sel[dev_id];
    size_t& cur_step = hw_step[dev_id];
    bool& np_set = hw_np[dev_id];
    std::lock_guard<std::mutex> lk(mu);
    cudaError_t e = cudaSuccess;
    if (h->smem_rows > cur_rows[h->NL]) {
      switch (h->NL) {
        case 1: e = cudaFuncSetAttribute(k_score_rows<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_rows); break;
        case 2: e = cudaFuncSetAttribute(k_score_rows<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_rows); break;
        default: e = cudaFuncSetAttribute(k_score_rows<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_rows); break;
      }
      if (e == cudaSuccess) cur_rows[h->NL] = h->smem_rows;
    }
    if (e == cudaSuccess && h->fused && h->smem_step > cur_step) {
      const int sm = (int)h->smem_step;
      e = cudaFuncSetAttribute(k_step<false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<false, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
      if (e == cudaSuccess) cur_step = h->smem_step;
    }
    if (e == cudaSuccess && h->cluster) {
      if (!np_set) {
        e = cudaFuncSetAttribute(k_step<false, true, false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<true, true, false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<false, true, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_step<true, true, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        np_set = (e == cudaSuccess);
      }
      // can at least one cluster of `beam` CTAs be resident?  otherwise fall back to the ticket variant
      int ncl_max = 0;
      cudaLaunchConfig_t qc = {};
      qc.gridDim = dim3(d.W, d.B);
      qc.blockDim = dim3(kStepThreads);
      qc.dynamicSmemBytes = h->smem_step;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = d.W;
      qa[0].val.clusterDim.y = 1;
      qa[0].val.clusterDim.z = 1;
      qc.attrs = qa;
      qc.numAttrs = 1;
      // clusters only pay while every utterance's cluster is resident at once (one wave): with more utterances
      // than that, whole-cluster scheduling leaves CTA slots idle (measured: 113 vs 94 us per launch at B = 512)
      if (e != cudaSuccess || cudaOccupancyMaxActiveClusters(&ncl_max, k_step<false, true, false>, &qc) != cudaSuccess ||
          ncl_max < d.B) {
        cudaGetLastError();
        e = cudaSuccess;
        h->cluster = 0;
      }
    }
    if (e == cudaSuccess) {
      // the SEQ instance (replay positions) pays only while the whole grid is one wave: its two halves run one
      // after the other in a CTA, which is hidden by the previous grid's tail at B = 64 (17.9 vs 19.6 us) but
      // costs slot time once CTAs queue for slots (109.9 vs 92.3 us per launch at B = 512)
      int per_sm = 0, dev = 0, sms = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_step<false, false, true>, kStepThreads, h->smem_step) !=
          cudaSuccess) {
        cudaGetLastError();
        per_sm = 0;
      }
      h->seq_ok = ((long long)per_sm * sms >= (long long)d.B * d.W) ? 1 : 0;
    }
    if (e == cudaSuccess && h->smem_sel > cur_sel) {
      e = cudaFuncSetAttribute(k_select, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_sel);
      if (e == cudaSuccess) cur_sel = h->smem_sel;
    }
    if (e != cudaSuccess) {
      lgasr_destroy(h);
      return fail(LGASR_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
  }
  // ---- k_step2 geometry: 6 scan rows of Tp + 4 floats, arena = pending keys | candidate rows | selection scratch ----
  {
    // k_step2 serves the term matrices k_step cannot read (bf16 storage, per-row normalisers: lgasr_step_ex) and the
    // plain fp32 launches it is faster on: a STREAMED row set (more than one chunk: two matrices at V = 5000, the
    // OWSM vocabulary) in a grid of several waves - its row stream runs on every warp with the rows prefetched into L2
    // (measured, profiles/r02x_shape_probe.txt: 123 vs 144 us per replay launch at configs[4]'s B = 512 with two
    // matrices, 73 vs 95 us on the OWSM shape at B = 128; one-chunk rows and one-wave grids stay with k_step: 18.1 vs
    // 19.2 us at configs[1]).  LGASR_V2=1 / 0 forces it on / off for every plain launch.
    const char* v2 = getenv("LGASR_V2");
    h->v2_all = (v2 && v2[0] == '1') ? 1 : 0;
    if (!(v2 && (v2[0] == '0' || v2[0] == '1')) && h->fused && h->sl.nchunks > 1 && !h->seq_ok) h->v2_all = 1;
    h->v2 = h->fused ? 1 : 0;
    if (h->v2) {
      StepLaunch& sl = h->sl2;
      sl = h->sl;
      const long long selw = utt_select_smem_floats(d.W, d.R, d.NS);
      // threads per CTA: the glue is paid per warp, so fewer, fatter warps issue fewer instructions per hypothesis
      {
        const char* nt = getenv("LGASR_NTH");
        h->nth2 = 128;  // measured: 128 threads per CTA beat 256 and 64 on the stepping launches (profiles/r02e)
        if (nt && (atoi(nt) == 128 || atoi(nt) == 256)) h->nth2 = atoi(nt);
      }
      // dense instance: a grid of several waves over streamed rows is bound by how many CTAs an SM holds (measured at
      // configs[4]'s B = 512 with two matrices, profiles/r02y_occupancy_probe.txt: 143 / 133 / 125 us per stepping launch
      // with 5 / 6 / 8 CTAs per SM): 64 registers per thread and the candidate rows in two groups (12 KB arena)
      int sms2 = 0, dev2 = 0;
      cudaGetDevice(&dev2);
      cudaDeviceGetAttribute(&sms2, cudaDevAttrMultiProcessorCount, dev2);
      const int V4b = (d.V + 3) & ~3;
      bool dense = h->nth2 == 128 && d.mode == kModePreBeam && (long long)(d.K > 0 ? d.K : 1) * V4b > 6144 &&
                   (long long)d.B * d.W > 6LL * sms2;
      if (const char* de = getenv("LGASR_DENSE")) dense = h->nth2 == 128 && de[0] == '1';
      h->mb2 = h->nth2 == 256 ? 5 : (dense ? 8 : 6);
      long long arena = (long long)n_sel * d.Tp;
      const long long acap = dense ? 3072 : 6144;
      if (arena > acap) arena = acap;
      if (arena < d.Tp) arena = d.Tp;
      if (d.mode != kModePreBeam) arena = 0;
      if (arena < 2 * kPend2) arena = 2 * kPend2;
      if (selw > arena) arena = selw;
      arena = (arena + 3) & ~3LL;
      sl.arena_floats = (int)arena;
      int rpg = (int)(arena / d.Tp);
      if (rpg < 1) rpg = 1;
      if (rpg > n_sel) rpg = n_sel;
      sl.rows_per_group = rpg;
      sl.kk_full = (n_sel + kStepWarps - 1) / kStepWarps;
      const char* ncl = getenv("LGASR_NO_CLUSTER");
      h->cluster2 = d.W >= 2 && d.W <= 16 && !(ncl && ncl[0] == '1');
      const size_t base = sizeof(Step2Smem) + (size_t)6 * (d.Tp + 4) * sizeof(float) + (size_t)arena * sizeof(float);
      h->smem_step2 = base + (h->cluster2 ? (size_t)selw * sizeof(float) : 0);
      if (h->smem_step2 > 220 * 1024) {
        lgasr_destroy(h);
        return fail(LGASR_EINVAL, "shared memory need too large for the fused step (T=%d beam=%d)", d.T, d.W);
      }
      // the attributes are per device: set them for every handle (cheap) instead of tracking high-water marks
      cudaError_t e = cudaSuccess;
      const void* fns[8];
      step2_instances(h->nth2, h->mb2, fns);
      static size_t hw2[64][3] = {};
      int dev_id = 0;
      cudaGetDevice(&dev_id);
      dev_id = dev_id < 0 || dev_id >= 64 ? 0 : dev_id;
      size_t& cur2 = hw2[dev_id][h->nth2 == 256 ? 0 : h->mb2 == 6 ? 1 : 2];
      // (the cluster instances get the attribute with the selection buffer included; a later fall-back to single CTAs
      // launches with less, which is fine)
      if (h->smem_step2 > cur2) {
        for (int i = 0; i < 8 && e == cudaSuccess; ++i)
          e = cudaFuncSetAttribute(fns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_step2);
        if (e == cudaSuccess) cur2 = h->smem_step2;
      }
      if (e == cudaSuccess && h->cluster2) {
        for (int i = 4; i < 8 && e == cudaSuccess; ++i)
          e = cudaFuncSetAttribute(fns[i], cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        int ncl_max = 0;
        cudaLaunchConfig_t qc = {};
        qc.gridDim = dim3(d.W, d.B);
        qc.blockDim = dim3(h->nth2);
        qc.dynamicSmemBytes = h->smem_step2;
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeClusterDimension;
        qa[0].val.clusterDim.x = d.W;
        qa[0].val.clusterDim.y = 1;
        qa[0].val.clusterDim.z = 1;
        qc.attrs = qa;
        qc.numAttrs = 1;
        // clusters only pay while every utterance's cluster is resident at once (one wave)
        if (e != cudaSuccess || cudaOccupancyMaxActiveClusters(&ncl_max, fns[5], &qc) != cudaSuccess || ncl_max < d.B) {
          cudaGetLastError();
          e = cudaSuccess;
          h->cluster2 = 0;
          h->smem_step2 = base;
        }
      }
      if (e != cudaSuccess) {
        lgasr_destroy(h);
        return fail(LGASR_ECUDA, "cudaFuncSetAttribute(k_step2): %s", cudaGetErrorString(e));
      }
    }
  }
  *out = h;
  return LGASR_OK;
}

extern "C" int lgasr_destroy(lgasr_handle* h) {
  if (!h) return LGASR_OK;
  for (void* p : h->allocs) cudaFree(p);
  for (cudaEvent_t e : h->prof_ev) cudaEventDestroy(e);
  if (h->rg_exec) cudaGraphExecDestroy(h->rg_exec);
  if (h->rg_stream) {
    cudaStreamDestroy(h->rg_stream);
    cudaEventDestroy(h->rg_ev[0]);
    cudaEventDestroy(h->rg_ev[1]);
  }
  if (h->h_pin) cudaFreeHost(h->h_pin);
  if (h->h_xlen_pin) cudaFreeHost(h->h_xlen_pin);
  if (h->ev_xlen) cudaEventDestroy(h->ev_xlen);
  if (h->h_begin_pin) cudaFreeHost(h->h_begin_pin);
  if (h->ev_begin) cudaEventDestroy(h->ev_begin);
  if (h->d_logp_stage) cudaFree(h->d_logp_stage);
  for (int s = 0; s < 2; ++s) {
    for (int k = 0; k < LGASR_MAX_TERMS; ++k)
      if (h->term_stage[s][k]) cudaFree(h->term_stage[s][k]);
    if (h->ev_copy[s]) cudaEventDestroy(h->ev_copy[s]);
    if (h->ev_comp[s]) cudaEventDestroy(h->ev_comp[s]);
  }
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  delete h;
  return LGASR_OK;
}

extern "C" int lgasr_set_posteriors(lgasr_handle* h, const float* d_logp, const int32_t* h_xlens, void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  cudaStream_t st = (cudaStream_t)stream;
  Dev& d = h->d;
  for (int b = 0; b < d.B; ++b) {
    int l = h_xlens ? h_xlens[b] : d.T;
    if (l < 1 || l > d.T) return fail(LGASR_EINVAL, "xlens[%d]=%d outside [1,%d]", b, l, d.T);
    h->xlens[b] = l;
  }
  // the lengths are staged in pinned memory owned by the handle: the call returns without synchronising the stream
  CK(cudaEventSynchronize(h->ev_xlen));  // the previous call's copy has left the staging buffer
  memcpy(h->h_xlen_pin, h->xlens.data(), sizeof(int) * d.B);
  CK(cudaMemcpyAsync(const_cast<int*>(d.xlen), h->h_xlen_pin, sizeof(int) * d.B, cudaMemcpyHostToDevice, st));
  CK(cudaEventRecord(h->ev_xlen, st));
  if (d.use_ctc) {
    if (!d_logp) return fail(LGASR_EINVAL, "null posteriors");
    dim3 grid((d.V + 63) / 64, (d.Tp + 63) / 64, d.B), block(32, 8);
    k_transpose_pad<<<grid, block, 0, st>>>(d_logp, const_cast<float*>(d.xT), d.xlen, d.T, d.Tp, d.V, d.blank, 0);
    LAUNCH_CK();
  }
  h->have_post = 1;
  h->begun = 0;
  return LGASR_OK;
}

extern "C" int lgasr_set_posteriors_host(lgasr_handle* h, const float* h_logp, const int32_t* h_xlens, void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  Dev& d = h->d;
  cudaStream_t st = (cudaStream_t)stream;
  if (d.use_ctc) {
    if (!h_logp) return fail(LGASR_EINVAL, "null posteriors");
    size_t n = (size_t)d.B * d.T * d.V;
    if (h->logp_stage_elems < n) {
      if (h->d_logp_stage) cudaFree(h->d_logp_stage);
      h->d_logp_stage = nullptr;
      h->logp_stage_elems = 0;
      if (cudaMalloc(&h->d_logp_stage, n * sizeof(float)) != cudaSuccess) return fail(LGASR_ENOMEM, "stage alloc");
      h->logp_stage_elems = n;
    }
    CK(cudaMemcpyAsync(h->d_logp_stage, h_logp, n * sizeof(float), cudaMemcpyHostToDevice, st));
  }
  return lgasr_set_posteriors(h, h->d_logp_stage, h_xlens, stream);
}

extern "C" int lgasr_begin(lgasr_handle* h, const int32_t* h_maxlen, const int32_t* h_minlen, const int64_t* h_primer,
                           void* stream) {
  if (!h || !h_maxlen || !h_minlen || !h_primer) return fail(LGASR_EINVAL, "null argument");
  if (!h->have_post) return fail(LGASR_ESTATE, "lgasr_begin before lgasr_set_posteriors");
  cudaStream_t st = (cudaStream_t)stream;
  Dev& d = h->d;
  for (int b = 0; b < d.B; ++b)
    if (h_maxlen[b] > d.maxlen_cap) return fail(LGASR_EINVAL, "maxlen[%d]=%d > cap %d", b, h_maxlen[b], d.maxlen_cap);
  h->primer.assign(h_primer, h_primer + h->cfg.primer_len);
  // the host arrays are staged in pinned memory owned by the handle, so the call does not have to
  // synchronise the stream before returning
  const int P = h->cfg.primer_len;
  if (!h->h_begin_pin) {
    if (cudaMallocHost(&h->h_begin_pin, sizeof(int) * 2 * d.B + sizeof(long long) * (P + 1)) != cudaSuccess)
      return fail(LGASR_ENOMEM, "cudaMallocHost");
    CK(cudaEventCreateWithFlags(&h->ev_begin, cudaEventDisableTiming));
  } else {
    CK(cudaEventSynchronize(h->ev_begin));  // the previous begin's copies have left the staging buffer
  }
  int* pin_i = h->h_begin_pin;
  long long* pin_p = reinterpret_cast<long long*>(pin_i + 2 * d.B + ((2 * d.B) & 1));
  memcpy(pin_i, h_maxlen, sizeof(int) * d.B);
  memcpy(pin_i + d.B, h_minlen, sizeof(int) * d.B);
  memcpy(pin_p, h_primer, sizeof(long long) * P);
  CK(cudaMemcpyAsync(const_cast<int*>(d.maxlen), pin_i, sizeof(int) * d.B, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(const_cast<int*>(d.minlen), pin_i + d.B, sizeof(int) * d.B, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(h->d_primer, pin_p, sizeof(long long) * P, cudaMemcpyHostToDevice, st));
  CK(cudaEventRecord(h->ev_begin, st));
  if (d.use_ctc) {
    const long long n = (long long)d.B * d.W * 2 * d.Tp;
    k_fill<<<592, 256, 0, st>>>(d.r[0], n, LGASR_LOGZERO);
    LAUNCH_CK();
    k_fill<<<592, 256, 0, st>>>(d.r[1], n, LGASR_LOGZERO);
    LAUNCH_CK();
    k_ctc_init_state<<<d.B, 128, d.Tp * sizeof(float), st>>>(d.xT, d.r[0], d.T, d.Tp, d.V, d.W, d.blank, d.W);
    LAUNCH_CK();
  }
  CK(cudaMemsetAsync(d.status, 0, sizeof(int), st));
  k_begin<<<d.B, 128, 0, st>>>(d, h->d_primer, h->cfg.primer_len, 0);
  LAUNCH_CK();
  h->pp = 0;
  h->rp = 0;
  h->step = 0;
  h->begun = 1;
  h->refilled = 0;
  h->cache_valid = false;
  return LGASR_OK;
}

extern "C" int lgasr_refill(lgasr_handle* h, int utt, const float* d_logp, int xlen, int maxlen, int minlen, void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  if (!h->begun) return fail(LGASR_ESTATE, "lgasr_refill before lgasr_begin");
  Dev& d = h->d;
  if (utt < 0 || utt >= d.B) return fail(LGASR_EINVAL, "utt %d out of range", utt);
  if (xlen < 1 || xlen > d.T) return fail(LGASR_EINVAL, "xlen %d outside [1,%d]", xlen, d.T);
  if (maxlen > d.maxlen_cap) return fail(LGASR_EINVAL, "maxlen %d > cap %d", maxlen, d.maxlen_cap);
  cudaStream_t st = (cudaStream_t)stream;
  h->xlens[utt] = xlen;
  k_set_utt<<<1, 1, 0, st>>>(d, utt, xlen, maxlen, minlen);
  LAUNCH_CK();
  if (d.use_ctc) {
    if (!d_logp) return fail(LGASR_EINVAL, "null posteriors");
    float* xT_u = const_cast<float*>(d.xT) + (size_t)utt * d.V * d.Tp;
    dim3 grid((d.V + 63) / 64, (d.Tp + 63) / 64, 1), block(32, 8);
    k_transpose_pad<<<grid, block, 0, st>>>(d_logp, xT_u, nullptr, d.T, d.Tp, d.V, d.blank, xlen);
    LAUNCH_CK();
    const long long n = (long long)d.W * 2 * d.Tp;
    for (int p = 0; p < 2; ++p) {
      k_fill<<<8, 256, 0, st>>>(d.r[p] + (size_t)utt * n, n, LGASR_LOGZERO);
      LAUNCH_CK();
    }
    // the initial state goes where the next position reads its parents
    k_ctc_init_state<<<1, 128, d.Tp * sizeof(float), st>>>(xT_u, d.r[h->rp] + (size_t)utt * n, d.T, d.Tp, 0, d.W, d.blank, d.W);
    LAUNCH_CK();
  }
  k_begin<<<1, 128, 0, st>>>(d, h->d_primer, h->cfg.primer_len, utt);
  LAUNCH_CK();
  h->refilled = 1;
  h->cache_valid = false;
  return LGASR_OK;
}

static int fill_terms(lgasr_handle* h, Terms& tm, const float* const* ptr, const float* cst, const float* w, int64_t ld,
                      const int32_t* dtype = nullptr, const float* const* norm = nullptr) {
  Dev& d = h->d;
  memset(&tm, 0, sizeof(tm));
  tm.K = d.K;
  tm.ld = ld;
  for (int k = 0; k < d.K; ++k) {
    tm.ptr[k] = ptr ? ptr[k] : nullptr;
    if (tm.ptr[k] && dtype) {
      if (dtype[k] != LGASR_F32 && dtype[k] != LGASR_BF16) return fail(LGASR_EINVAL, "term %d: unknown dtype %d", k, dtype[k]);
      tm.bf16[k] = dtype[k] == LGASR_BF16 ? 1 : 0;
    }
    tm.norm[k] = (tm.ptr[k] && norm) ? norm[k] : nullptr;
    tm.cst[k] = cst ? cst[k] : 0.f;
    tm.w[k] = w ? w[k] : 0.f;
    tm.slot[k] = tm.nptr;
    if (tm.ptr[k]) ++tm.nptr;
  }
  if (d.key >= 0 && !tm.ptr[d.key]) return fail(LGASR_EINVAL, "pre_beam_key term %d has no matrix", d.key);
  if (ld < d.V && d.K > 0) return fail(LGASR_EINVAL, "ld %lld < V %d", (long long)ld, d.V);
  return LGASR_OK;
}

static int step_impl(lgasr_handle* h, const Terms& tm, cudaStream_t st, bool pdl = false, bool flags = false) {
  Dev& d = h->d;
  if (!h->refilled && h->step >= d.maxlen_cap)
    return fail(LGASR_ESTATE, "step %d beyond max_len_cap %d", h->step, d.maxlen_cap);
  dim3 grid(d.W, d.B);
  cudaEvent_t pe[3] = {nullptr, nullptr, nullptr};
  if (h->profiling) {
    for (int i = 0; i < 3; ++i) {
      CK(cudaEventCreate(&pe[i]));
      h->prof_ev.push_back(pe[i]);
    }
    CK(cudaEventRecord(pe[0], st));
  }
  // k_step2 reads the rows with 16-byte loads: aligned rows whose length is a multiple of 4 entries (anything else,
  // fp32 only, takes k_step, which can stage unaligned rows with plain loads)
  bool al = (tm.ld % 4 == 0) && (d.V % 4 == 0), plain = true;
  for (int k = 0; k < tm.K; ++k)
    if (tm.ptr[k]) {
      al = al && ((reinterpret_cast<uintptr_t>(tm.ptr[k]) & 15) == 0);
      plain = plain && !tm.bf16[k] && !tm.norm[k];
    }
  if (!plain && !(h->v2 && al))
    return fail(LGASR_EINVAL, "bf16 / normalised term matrices need 16-byte aligned rows, V %% 4 == 0 and the fused step");
  if (!plain && d.key >= 0 && (tm.bf16[d.key] || tm.norm[d.key]))
    return fail(LGASR_EINVAL, "the pre-beam key term must be a plain fp32 log-probability matrix");
  if (h->v2 && al && (h->v2_all || !plain)) {
    StepLaunch sl = h->sl2;
    sl.tma_rows = 1;
    sl.row_path = (d.key >= 0) ? 1 : (plain && tm.K == 1 && tm.nptr == 1) ? 1 : (plain && tm.K == 2 && tm.nptr == 2) ? 2 : 0;
    sl.pdl = pdl ? 1 : 0;
    sl.flags = 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kStepThreads);
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (pdl) {
      at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[na].val.programmaticStreamSerializationAllowed = 1;
      ++na;
    }
    // replay positions (programmatic launches) run as single CTAs with the ticket hand-off: clusters of `beam` CTAs
    // only become resident when `beam` slots of one GPC are free at once, which delays the early start they live on
    const char* pcl = getenv("LGASR_PDL_CLUSTER");
    const bool use_cluster = h->cluster2 && (!pdl || (pcl && pcl[0] == '1'));
    if (use_cluster) {
      at[na].id = cudaLaunchAttributeClusterDimension;
      at[na].val.clusterDim.x = d.W;
      at[na].val.clusterDim.y = 1;
      at[na].val.clusterDim.z = 1;
      ++na;
    }
    cfg.attrs = at;
    cfg.numAttrs = na;
    // without the cluster the selection scratch aliases the arena: the extra buffer is not needed
    cfg.dynamicSmemBytes = h->smem_step2;
    const bool trc = d.trace != nullptr && sl.row_path == 1;  // the stamped instance exists for the one-matrix path
    const void* fns[8];
    step2_instances(h->nth2, h->mb2, fns);
    const void* kern = fns[(use_cluster ? 4 : 0) + (trc ? 3 : sl.row_path)];
    cfg.blockDim = dim3(h->nth2);
    void* args[5] = {(void*)&d, (void*)&tm, (void*)&sl, (void*)&h->pp, (void*)&h->rp};
    CK(cudaLaunchKernelExC(&cfg, kern, args));
    LAUNCH_CK();
    if (h->profiling) {
      CK(cudaEventRecord(pe[1], st));
      CK(cudaEventRecord(pe[2], st));
    }
  } else if (h->fused) {
    StepLaunch sl = h->sl;
    // TMA bulk copies need 16-byte aligned rows whose chunks are multiples of 16 bytes
    bool al = (tm.nptr > 0) && (tm.ld % 4 == 0) && (d.V % 4 == 0);
    for (int k = 0; k < tm.K; ++k)
      if (tm.ptr[k]) al = al && ((reinterpret_cast<uintptr_t>(tm.ptr[k]) & 15) == 0);
    sl.tma_rows = al ? 1 : 0;
    // programmatic dependent launch (only between consecutive k_step launches of a replay, where the
    // matrices are inputs that no earlier kernel of the stream writes): see the prologue of k_step
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kStepThreads);
    cfg.dynamicSmemBytes = h->smem_step;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    int na = 0;
    // Stepping launches carry the attribute too: the CTAs' prologue (mbarrier set-up, parameter and metadata address
    // arithmetic) then overlaps the tail of the previous kernel of the stream when that kernel lets its dependents go
    // early (k_step, k_log_softmax_rows and k_score_filter do); everything that reads memory comes after
    // griddepcontrol.wait, and after a kernel that never triggers this is an ordinary launch.  22.0 -> 20.0 us per
    // launch at B = 64 (profiles/r02af_step_pdl_probe.txt).  LGASR_STEP_PDL=0 switches it off.
    const char* spd = getenv("LGASR_STEP_PDL");
    if (pdl || !(spd && spd[0] == '0')) {
      at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[na].val.programmaticStreamSerializationAllowed = 1;
      ++na;
    }
    // SEQ variant (selection before griddepcontrol.wait, four CTC warps after it) for programmatic launches.
    // It runs as single CTAs with the ticket hand-off: clusters of `beam` CTAs only become resident when ten
    // slots of one GPC are free at once, which delays the early start the variant lives on (measured: 20.2 us
    // per launch with clusters, 17.9 us without).
    const char* nseq = getenv("LGASR_NO_SEQ");
    const bool seq = pdl && h->seq_ok && !(nseq && nseq[0] == '1');
    const bool use_cluster = h->cluster && !seq;
    if (use_cluster) {
      at[na].id = cudaLaunchAttributeClusterDimension;
      at[na].val.clusterDim.x = d.W;
      at[na].val.clusterDim.y = 1;
      at[na].val.clusterDim.z = 1;
      ++na;
    }
    cfg.attrs = at;
    cfg.numAttrs = na;
    sl.pdl = pdl ? 1 : 0;
    sl.flags = flags ? 1 : 0;
    auto kern = k_step<false, false, false>;
    if (use_cluster) {
      if (d.trace) kern = seq ? k_step<true, true, true> : k_step<true, true, false>;
      else kern = seq ? k_step<false, true, true> : k_step<false, true, false>;
    } else {
      if (d.trace) kern = seq ? k_step<true, false, true> : k_step<true, false, false>;
      else kern = seq ? k_step<false, false, true> : k_step<false, false, false>;
    }
    CK(cudaLaunchKernelEx(&cfg, kern, d, tm, sl, h->pp, h->rp, h->step));
    LAUNCH_CK();
    if (h->profiling) {
      CK(cudaEventRecord(pe[1], st));
      CK(cudaEventRecord(pe[2], st));
    }
  } else {
    switch (h->NL) {
      case 1: k_score_rows<1><<<grid, kRowThreads, h->smem_rows, st>>>(d, tm, h->pp, h->rp); break;
      case 2: k_score_rows<2><<<grid, kRowThreads, h->smem_rows, st>>>(d, tm, h->pp, h->rp); break;
      default: k_score_rows<4><<<grid, kRowThreads, h->smem_rows, st>>>(d, tm, h->pp, h->rp); break;
    }
    LAUNCH_CK();
    if (h->profiling) CK(cudaEventRecord(pe[1], st));
    k_select<<<d.B, kSelThreads, h->smem_sel, st>>>(d, tm, h->pp, h->step);
    LAUNCH_CK();
    if (h->profiling) CK(cudaEventRecord(pe[2], st));
  }
  h->pp ^= 1;
  h->rp ^= 1;
  h->step += 1;
  h->cache_valid = false;
  return LGASR_OK;
}

extern "C" int lgasr_step(lgasr_handle* h, const float* const* h_term_ptr, const float* h_term_const,
                          const float* h_term_weight, int64_t ld, void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  if (!h->begun) return fail(LGASR_ESTATE, "lgasr_step before lgasr_begin");
  Terms tm;
  TRY(fill_terms(h, tm, h_term_ptr, h_term_const, h_term_weight, ld));
  return step_impl(h, tm, (cudaStream_t)stream);
}

extern "C" int lgasr_step_ex(lgasr_handle* h, const void* const* h_term_ptr, const int32_t* h_term_dtype,
                             const float* const* h_term_norm, const float* h_term_const, const float* h_term_weight, int64_t ld,
                             void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  if (!h->begun) return fail(LGASR_ESTATE, "lgasr_step_ex before lgasr_begin");
  Terms tm;
  TRY(fill_terms(h, tm, reinterpret_cast<const float* const*>(h_term_ptr), h_term_const, h_term_weight, ld, h_term_dtype,
                 h_term_norm));
  return step_impl(h, tm, (cudaStream_t)stream);
}

extern "C" int lgasr_launch_info(lgasr_handle* h, int32_t* cluster, int32_t* smem_bytes, int32_t* chunk_floats, int32_t* stages) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  if (cluster) *cluster = h->v2_all ? h->cluster2 : (h->fused ? h->cluster : 0);
  if (smem_bytes) *smem_bytes = (int32_t)(h->v2_all ? h->smem_step2 : (h->fused ? h->smem_step : h->smem_rows));
  if (chunk_floats) *chunk_floats = h->sl.CH;
  if (stages) *stages = h->sl.NST;
  return LGASR_OK;
}

extern "C" int lgasr_set_profiling(lgasr_handle* h, int on) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  for (cudaEvent_t e : h->prof_ev) cudaEventDestroy(e);
  h->prof_ev.clear();
  h->profiling = on ? 1 : 0;
  return LGASR_OK;
}

extern "C" int lgasr_get_profile(lgasr_handle* h, float* ms_score_rows, float* ms_select, int32_t* n_steps) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  double a = 0.0, b = 0.0;
  const size_t n = h->prof_ev.size() / 3;
  for (size_t i = 0; i < n; ++i) {
    float t0 = 0.f, t1 = 0.f;
    CK(cudaEventSynchronize(h->prof_ev[3 * i + 2]));
    CK(cudaEventElapsedTime(&t0, h->prof_ev[3 * i], h->prof_ev[3 * i + 1]));
    CK(cudaEventElapsedTime(&t1, h->prof_ev[3 * i + 1], h->prof_ev[3 * i + 2]));
    a += t0;
    b += t1;
  }
  if (ms_score_rows) *ms_score_rows = (float)a;
  if (ms_select) *ms_select = (float)b;
  if (n_steps) *n_steps = (int)n;
  return LGASR_OK;
}

extern "C" int lgasr_debug_trace(lgasr_handle* h, int64_t* h_out, int64_t n) {
  if (!h || !h_out) return fail(LGASR_EINVAL, "null argument");
  if (!h->d.trace) return fail(LGASR_ESTATE, "tracing is off (set LGASR_TRACE=1 before lgasr_create)");
  const int64_t have = (int64_t)h->d.B * h->d.W * 32;
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(h_out, h->d.trace, sizeof(long long) * (size_t)(n < have ? n : have), cudaMemcpyDeviceToHost));
  return LGASR_OK;
}

extern "C" int lgasr_view(lgasr_handle* h, int which, void** d_ptr, int64_t* ld) {
  if (!h || !d_ptr) return fail(LGASR_EINVAL, "null argument");
  Dev& d = h->d;
  int64_t l = 0;
  switch (which) {
    case 0: *d_ptr = d.yseq[h->pp]; l = d.Lcap; break;
    case 1: *d_ptr = d.n_run[h->pp]; l = 1; break;
    case 2: *d_ptr = d.step_parent; l = d.W; break;
    case 3: *d_ptr = d.step_tok; l = d.W; break;
    case 4: *d_ptr = d.done[h->pp]; l = 1; break;
    case 5: *d_ptr = d.score[h->pp]; l = d.W; break;
    case 6: *d_ptr = d.cur_len[h->pp]; l = 1; break;
    default: return fail(LGASR_EINVAL, "unknown view %d", which);
  }
  if (ld) *ld = l;
  return LGASR_OK;
}

extern "C" int lgasr_poll(lgasr_handle* h, int32_t* h_n_running, int32_t* h_done, int32_t* h_parent, int32_t* h_token,
                          int32_t* n_active, void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  cudaStream_t st = (cudaStream_t)stream;
  Dev& d = h->d;
  const int BW = d.B * d.W;
  int* pin = h->h_pin;
  CK(cudaMemcpyAsync(pin, d.n_run[h->pp], sizeof(int) * d.B, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(pin + d.B, d.done[h->pp], sizeof(int) * d.B, cudaMemcpyDeviceToHost, st));
  if (h_parent) CK(cudaMemcpyAsync(pin + 2 * d.B, d.step_parent, sizeof(int) * BW, cudaMemcpyDeviceToHost, st));
  if (h_token) CK(cudaMemcpyAsync(pin + 2 * d.B + BW, d.step_tok, sizeof(int) * BW, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(pin + 2 * d.B + 2 * BW, d.status, sizeof(int), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (pin[2 * d.B + 2 * BW] & 1) return fail(LGASR_ESTATE, "ended-hypotheses pool overflow (Ecap = beam * max_len_cap = %d)", d.Ecap);
  int act = 0;
  for (int b = 0; b < d.B; ++b) {
    if (h_n_running) h_n_running[b] = pin[b];
    if (h_done) h_done[b] = pin[d.B + b];
    if (!pin[d.B + b]) ++act;
  }
  if (h_parent) memcpy(h_parent, pin + 2 * d.B, sizeof(int) * BW);
  if (h_token) memcpy(h_token, pin + 2 * d.B + BW, sizeof(int) * BW);
  if (n_active) *n_active = act;
  return LGASR_OK;
}

static int refresh_cache(lgasr_handle* h, cudaStream_t st) {
  Dev& d = h->d;
  h->c_n_ended.resize(d.B);
  int status = 0;
  CK(cudaMemcpyAsync(h->c_n_ended.data(), d.n_ended, sizeof(int) * d.B, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&status, d.status, sizeof(int), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (status & 1) return fail(LGASR_ESTATE, "ended-hypotheses pool overflow (Ecap = beam * max_len_cap = %d)", d.Ecap);
  int mx = 0;
  for (int b = 0; b < d.B; ++b) mx = h->c_n_ended[b] > mx ? h->c_n_ended[b] : mx;
  const int NS = d.NS > 0 ? d.NS : 1;
  const int steps = (h->refilled || h->step > d.maxlen_cap) ? d.maxlen_cap : h->step;  // trellis rows in use
  h->cache_max_n = mx;
  h->cache_steps = steps;
  h->c_score.resize((size_t)d.B * (mx ? mx : 1));
  h->c_meta.resize((size_t)d.B * (mx ? mx : 1));
  h->c_scores.resize((size_t)d.B * (mx ? mx : 1) * NS);
  h->c_bp_par.resize((size_t)d.B * (steps ? steps : 1) * d.W);
  h->c_bp_tok.resize((size_t)d.B * (steps ? steps : 1) * d.W);
  if (mx > 0) {
    CK(cudaMemcpy2DAsync(h->c_score.data(), sizeof(float) * mx, d.end_score, sizeof(float) * d.Ecap, sizeof(float) * mx, d.B,
                         cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpy2DAsync(h->c_meta.data(), sizeof(int4) * mx, d.end_meta, sizeof(int4) * d.Ecap, sizeof(int4) * mx, d.B,
                         cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpy2DAsync(h->c_scores.data(), sizeof(float) * mx * NS, d.end_scores, sizeof(float) * d.Ecap * NS,
                         sizeof(float) * mx * NS, d.B, cudaMemcpyDeviceToHost, st));
  }
  if (steps > 0) {
    CK(cudaMemcpy2DAsync(h->c_bp_par.data(), sizeof(int) * steps * d.W, d.bp_par, sizeof(int) * d.maxlen_cap * d.W,
                         sizeof(int) * steps * d.W, d.B, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpy2DAsync(h->c_bp_tok.data(), sizeof(int) * steps * d.W, d.bp_tok, sizeof(int) * d.maxlen_cap * d.W,
                         sizeof(int) * steps * d.W, d.B, cudaMemcpyDeviceToHost, st));
  }
  CK(cudaStreamSynchronize(st));
  h->cache_valid = true;
  return LGASR_OK;
}

extern "C" int lgasr_gather_ended(lgasr_handle* h, int utt, int32_t* n_ended, int32_t* max_len, int64_t* h_tokens,
                                  int32_t* h_lens, float* h_score, float* h_scores, void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  Dev& d = h->d;
  if (utt < 0 || utt >= d.B) return fail(LGASR_EINVAL, "utt %d out of range", utt);
  if (!h->begun) return fail(LGASR_ESTATE, "gather before begin");
  if (!h->cache_valid) TRY(refresh_cache(h, (cudaStream_t)stream));
  const int n = h->c_n_ended[utt];
  const int mx = h->cache_max_n, steps = h->cache_steps;
  const int NS = d.NS > 0 ? d.NS : 1;
  int ml = 0;
  for (int e = 0; e < n; ++e) {
    int l = h->c_meta[(size_t)utt * mx + e].z;
    ml = l > ml ? l : ml;
  }
  if (n_ended) *n_ended = n;
  if (max_len) *max_len = ml;
  if (!h_tokens) return LGASR_OK;
  const int P = h->cfg.primer_len;
  for (int e = 0; e < n; ++e) {
    const int4 m = h->c_meta[(size_t)utt * mx + e];  // step, slot, len, extra_eos
    int64_t* row = h_tokens + (size_t)e * ml;
    for (int t = 0; t < ml; ++t) row[t] = -1;
    for (int t = 0; t < P; ++t) row[t] = h->primer[t];
    int slot = m.y;
    for (int s = m.x; s >= 0; --s) {
      if (slot < 0) return fail(LGASR_ESTATE, "broken back-pointer chain (utt %d, ended %d)", utt, e);
      const size_t o = ((size_t)utt * steps + s) * d.W + slot;
      row[P + s] = h->c_bp_tok[o];
      slot = h->c_bp_par[o];
    }
    if (m.w) row[P + m.x + 1] = d.eos;
    if (h_lens) h_lens[e] = m.z;
    if (h_score) h_score[e] = h->c_score[(size_t)utt * mx + e];
    if (h_scores)
      for (int k = 0; k < d.NS; ++k) h_scores[(size_t)e * d.NS + k] = h->c_scores[((size_t)utt * mx + e) * NS + k];
  }
  return LGASR_OK;
}

extern "C" int lgasr_gather_best(lgasr_handle* h, int normalize_length, int max_len, int64_t* h_tokens, int32_t* h_lens,
                                 float* h_score, void* stream) {
  if (!h || !h_tokens || !h_lens || !h_score) return fail(LGASR_EINVAL, "null argument");
  Dev& d = h->d;
  if (!h->begun) return fail(LGASR_ESTATE, "gather before begin");
  if (!h->cache_valid) TRY(refresh_cache(h, (cudaStream_t)stream));
  const int mx = h->cache_max_n, steps = h->cache_steps;
  const int P = h->cfg.primer_len;
  for (int utt = 0; utt < d.B; ++utt) {
    const int n = h->c_n_ended[utt];
    int64_t* row = h_tokens + (size_t)utt * max_len;
    for (int t = 0; t < max_len; ++t) row[t] = -1;
    int best = -1;
    double bkey = 0.0;
    for (int e = 0; e < n; ++e) {
      const int4 m = h->c_meta[(size_t)utt * mx + e];
      const float sc = h->c_score[(size_t)utt * mx + e];
      // beam_search.py:452-459: python floats (double) of the fp32 score, / (len(yseq) - 1) when normalising
      const double key = normalize_length ? (double)sc / (double)(m.z - 1) : (double)sc;
      if (best < 0 || key > bkey) {
        best = e;
        bkey = key;
      }
    }
    h_lens[utt] = 0;
    h_score[utt] = 0.f;
    if (best < 0) continue;
    const int4 m = h->c_meta[(size_t)utt * mx + best];  // step, slot, len, extra_eos
    if (m.z > max_len) return fail(LGASR_EINVAL, "max_len %d < hypothesis length %d", max_len, m.z);
    for (int t = 0; t < P; ++t) row[t] = h->primer[t];
    int slot = m.y;
    for (int s2 = m.x; s2 >= 0; --s2) {
      if (slot < 0) return fail(LGASR_ESTATE, "broken back-pointer chain (utt %d, ended %d)", utt, best);
      const size_t o = ((size_t)utt * steps + s2) * d.W + slot;
      row[P + s2] = h->c_bp_tok[o];
      slot = h->c_bp_par[o];
    }
    if (m.w) row[P + m.x + 1] = d.eos;
    h_lens[utt] = m.z;
    h_score[utt] = h->c_score[(size_t)utt * mx + best];
  }
  return LGASR_OK;
}

extern "C" int lgasr_decode_replay(lgasr_handle* h, int n_steps, const float* const* h_term_ptr, const float* h_term_const,
                                   const float* h_term_weight, int64_t ld, void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  if (!h->begun) return fail(LGASR_ESTATE, "replay before begin");
  const int K = h->d.K;
  cudaStream_t user_st = (cudaStream_t)stream;
  cudaStream_t st = user_st;
  const char* npdl = getenv("LGASR_NO_PDL");
  const bool pdl = !h->profiling && !(npdl && npdl[0] == '1');
  // experimental (off by default, +1 % at configs[1]): per-utterance release/acquire flags instead of the grid-wide
  // griddepcontrol.wait between the positions of a replay
  const char* fl = getenv("LGASR_FLAGS");
  const bool flags = pdl && fl && fl[0] == '1';
  auto run_steps = [&]() -> int {
    for (int i = 0; i < n_steps; ++i) {
      Terms tm;
      TRY(fill_terms(h, tm, h_term_ptr ? h_term_ptr + (size_t)i * K : nullptr, h_term_const, h_term_weight, ld));
      TRY(step_impl(h, tm, st, pdl && i > 0, flags));
    }
    return LGASR_OK;
  };
  const char* ng = getenv("LGASR_NO_GRAPH");
  if (h->profiling || n_steps < 2 || !h_term_ptr || (ng && ng[0] == '1')) return run_steps();
  if (!h->rg_stream) {
    CK(cudaStreamCreateWithFlags(&h->rg_stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&h->rg_ev[0], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->rg_ev[1], cudaEventDisableTiming));
  }
  st = h->rg_stream;
  // one launch per position and nothing on the host between them: the loop is launch-bound, so it is
  // captured once into a CUDA graph and replayed while the same buffers are decoded again
  std::vector<const float*> ptrs(h_term_ptr, h_term_ptr + (size_t)n_steps * K);
  std::vector<float> cst(K), w(K);
  for (int k = 0; k < K; ++k) {
    cst[k] = h_term_const ? h_term_const[k] : 0.f;
    w[k] = h_term_weight ? h_term_weight[k] : 0.f;
  }
  const bool hit = h->rg_exec && h->rg_steps == n_steps && h->rg_ld == (long long)ld && h->rg_pp == h->pp &&
                   h->rg_rp == h->rp && h->rg_step0 == h->step && h->rg_ptrs == ptrs && h->rg_cst == cst && h->rg_w == w;
  if (!hit) {
    if (h->step + n_steps > h->d.maxlen_cap) return fail(LGASR_ESTATE, "replay beyond max_len_cap");
    if (h->rg_exec) {
      cudaGraphExecDestroy(h->rg_exec);
      h->rg_exec = nullptr;
    }
    const int pp0 = h->pp, rp0 = h->rp, step0 = h->step;
    cudaGraph_t g = nullptr;
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    const int rc = run_steps();
    g_launches.fetch_sub(n_steps, std::memory_order_relaxed);  // captured, not launched
    const cudaError_t ee = cudaStreamEndCapture(st, &g);
    h->pp = pp0;
    h->rp = rp0;
    h->step = step0;
    if (rc) {
      if (g) cudaGraphDestroy(g);
      return rc;
    }
    if (ee != cudaSuccess) return fail(LGASR_ECUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(ee));
    if (const char* dot = getenv("LGASR_GRAPH_DOT")) cudaGraphDebugDotPrint(g, dot, cudaGraphDebugDotFlagsVerbose);
    const cudaError_t ei = cudaGraphInstantiate(&h->rg_exec, g, 0);
    cudaGraphDestroy(g);
    if (ei != cudaSuccess) {
      h->rg_exec = nullptr;
      return fail(LGASR_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(ei));
    }
    h->rg_ptrs.swap(ptrs);
    h->rg_cst = cst;
    h->rg_w = w;
    h->rg_ld = ld;
    h->rg_steps = n_steps;
    h->rg_pp = pp0;
    h->rg_rp = rp0;
    h->rg_step0 = step0;
  } else if (h->step + n_steps > h->d.maxlen_cap) {
    return fail(LGASR_ESTATE, "replay beyond max_len_cap");
  }
  CK(cudaEventRecord(h->rg_ev[0], user_st));
  CK(cudaStreamWaitEvent(st, h->rg_ev[0], 0));
  CK(cudaGraphLaunch(h->rg_exec, st));
  CK(cudaEventRecord(h->rg_ev[1], st));
  CK(cudaStreamWaitEvent(user_st, h->rg_ev[1], 0));
  g_launches.fetch_add(n_steps, std::memory_order_relaxed);
  if (n_steps & 1) {
    h->pp ^= 1;
    h->rp ^= 1;
  }
  h->step += n_steps;
  h->cache_valid = false;
  return LGASR_OK;
}

extern "C" int lgasr_decode_replay_host(lgasr_handle* h, const float* h_logp, const int32_t* h_xlens, const int32_t* h_maxlen,
                                        const int32_t* h_minlen, const int64_t* h_primer, int n_steps,
                                        const float* const* h_term_ptr, const float* h_term_const,
                                        const float* h_term_weight, int64_t ld, void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  Dev& d = h->d;
  cudaStream_t st = (cudaStream_t)stream;
  const int K = d.K;
  const size_t elems = (size_t)d.B * d.W * ld;
  if (!h->copy_stream) {
    CK(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int s = 0; s < 2; ++s) {
      CK(cudaEventCreateWithFlags(&h->ev_copy[s], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&h->ev_comp[s], cudaEventDisableTiming));
    }
  }
  if (h->term_stage_elems < elems) {
    for (int s = 0; s < 2; ++s)
      for (int k = 0; k < LGASR_MAX_TERMS; ++k) {
        if (h->term_stage[s][k]) cudaFree(h->term_stage[s][k]);
        h->term_stage[s][k] = nullptr;
      }
    h->term_stage_elems = 0;
    for (int s = 0; s < 2; ++s)
      for (int k = 0; k < K; ++k)
        if (cudaMalloc(&h->term_stage[s][k], elems * sizeof(float)) != cudaSuccess) return fail(LGASR_ENOMEM, "term staging");
    h->term_stage_elems = elems;
  }
  // the first matrices travel while the posteriors are transposed
  auto upload = [&](int i) -> int {
    const int s = i & 1;
    if (i >= 2) CK(cudaStreamWaitEvent(h->copy_stream, h->ev_comp[s], 0));
    for (int k = 0; k < K; ++k) {
      const float* src = h_term_ptr[(size_t)i * K + k];
      if (src) CK(cudaMemcpyAsync(h->term_stage[s][k], src, elems * sizeof(float), cudaMemcpyHostToDevice, h->copy_stream));
    }
    CK(cudaEventRecord(h->ev_copy[s], h->copy_stream));
    return LGASR_OK;
  };
  TRY(lgasr_set_posteriors_host(h, h_logp, h_xlens, stream));
  TRY(lgasr_begin(h, h_maxlen, h_minlen, h_primer, stream));
  if (n_steps > 0) TRY(upload(0));
  for (int i = 0; i < n_steps; ++i) {
    const int s = i & 1;
    if (i + 1 < n_steps) TRY(upload(i + 1));
    CK(cudaStreamWaitEvent(st, h->ev_copy[s], 0));
    const float* ptrs[LGASR_MAX_TERMS];
    for (int k = 0; k < K; ++k) ptrs[k] = h_term_ptr[(size_t)i * K + k] ? h->term_stage[s][k] : nullptr;
    Terms tm;
    TRY(fill_terms(h, tm, ptrs, h_term_const, h_term_weight, ld));
    TRY(step_impl(h, tm, st));
    CK(cudaEventRecord(h->ev_comp[s], st));
  }
  return LGASR_OK;
}

// ---- stand-alone scorer ABI ---------------------------------------------------------------------
extern "C" int lgasr_transpose_posteriors(const float* d_logp, int T, int V, int xlen, int blank, float* d_xT, int Tp,
                                          void* stream) {
  if (!d_logp || !d_xT || T < 1 || V < 1 || Tp < T) return fail(LGASR_EINVAL, "bad argument");
  dim3 grid((V + 63) / 64, (Tp + 63) / 64, 1), block(32, 8);
  k_transpose_pad<<<grid, block, 0, (cudaStream_t)stream>>>(d_logp, d_xT, nullptr, T, Tp, V, blank, xlen);
  LAUNCH_CK();
  return LGASR_OK;
}

extern "C" int lgasr_ctc_initial_state(const float* d_xT, int T, int Tp, int blank, float* d_r0, void* stream) {
  if (!d_xT || !d_r0 || T < 1 || Tp < T) return fail(LGASR_EINVAL, "bad argument");
  // V is only used for the per-utterance stride; one utterance here
  k_ctc_init_state<<<1, 128, Tp * sizeof(float), (cudaStream_t)stream>>>(d_xT, d_r0, T, Tp, 0, 1, blank, 1);
  LAUNCH_CK();
  return LGASR_OK;
}

extern "C" int lgasr_ctc_prefix_score(const float* d_xT, int T, int Tp, int V, int blank, int eos, int n,
                                      const float* d_r_prev, const float* d_s_prev, const int32_t* d_last, int olen,
                                      const int64_t* d_ids, int S, float* d_scores, float* d_log_psi, void* stream) {
  return lgasr_ctc_prefix_score_win(d_xT, T, Tp, V, blank, eos, n, d_r_prev, d_s_prev, d_last, olen, d_ids, S, d_scores, d_log_psi,
                                    0, 0, stream);
}

extern "C" int lgasr_ctc_prefix_score_win(const float* d_xT, int T, int Tp, int V, int blank, int eos, int n,
                                          const float* d_r_prev, const float* d_s_prev, const int32_t* d_last, int olen,
                                          const int64_t* d_ids, int S, float* d_scores, float* d_log_psi, int win_start,
                                          int win_end, void* stream) {
  if (win_start < 0 || win_end < 0 || (win_end > 0 && win_start >= win_end) || win_end > T) return fail(LGASR_EINVAL, "bad window");
  if (!d_xT || !d_r_prev || !d_last || !d_scores || !d_log_psi || n < 1 || T < 1 || Tp < T)
    return fail(LGASR_EINVAL, "bad argument");
  if (S > 0 && !d_ids) return fail(LGASR_EINVAL, "null ids");
  const size_t smem = (size_t)2 * Tp * sizeof(float);
  if (smem > 200 * 1024) return fail(LGASR_EINVAL, "T too large");
  {
    static std::mutex mu;
    static size_t hw[64] = {};  // per device (the attribute is per function and device)
    int dev_id = 0;
    cudaGetDevice(&dev_id);
    size_t& cur = hw[dev_id < 0 || dev_id >= 64 ? 0 : dev_id];
    std::lock_guard<std::mutex> lk(mu);
    if (smem > 48 * 1024 && smem > cur) {
      cudaError_t e = cudaFuncSetAttribute(k_ctc_prefix_score, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return fail(LGASR_ECUDA, "%s", cudaGetErrorString(e));
      cur = smem;
    }
  }
  k_ctc_prefix_score<<<n, 256, smem, (cudaStream_t)stream>>>(d_xT, T, Tp, V, blank, eos, d_r_prev, d_s_prev, d_last, olen,
                                                            reinterpret_cast<const long long*>(d_ids), S, d_scores, d_log_psi,
                                                            win_start, win_end);
  LAUNCH_CK();
  return LGASR_OK;
}

extern "C" int lgasr_ctc_select_state(const float* d_xT, int T, int Tp, int blank, int m, const float* d_r_prev,
                                      const int32_t* d_last, const int32_t* d_sel_hyp, const int32_t* d_sel_tok, int olen,
                                      float* d_r_new, void* stream) {
  return lgasr_ctc_select_state_win(d_xT, T, Tp, blank, m, d_r_prev, d_last, d_sel_hyp, d_sel_tok, olen, d_r_new, 0, 0, stream);
}

extern "C" int lgasr_ctc_select_state_win(const float* d_xT, int T, int Tp, int blank, int m, const float* d_r_prev,
                                          const int32_t* d_last, const int32_t* d_sel_hyp, const int32_t* d_sel_tok, int olen,
                                          float* d_r_new, int win_start, int win_end, void* stream) {
  if (win_start < 0 || win_end < 0 || (win_end > 0 && win_start >= win_end) || win_end > T) return fail(LGASR_EINVAL, "bad window");
  if (!d_xT || !d_r_prev || !d_last || !d_sel_hyp || !d_sel_tok || !d_r_new || m < 1 || T < 1 || Tp < T)
    return fail(LGASR_EINVAL, "bad argument");
  const size_t smem = (size_t)4 * Tp * sizeof(float);
  if (smem > 200 * 1024) return fail(LGASR_EINVAL, "T too large");
  {
    static std::mutex mu;
    static size_t hw[64] = {};  // per device (the attribute is per function and device)
    int dev_id = 0;
    cudaGetDevice(&dev_id);
    size_t& cur = hw[dev_id < 0 || dev_id >= 64 ? 0 : dev_id];
    std::lock_guard<std::mutex> lk(mu);
    if (smem > 48 * 1024 && smem > cur) {
      cudaError_t e = cudaFuncSetAttribute(k_ctc_select_state, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return fail(LGASR_ECUDA, "%s", cudaGetErrorString(e));
      cur = smem;
    }
  }
  k_ctc_select_state<<<m, 32, smem, (cudaStream_t)stream>>>(d_xT, T, Tp, blank, d_r_prev, d_last, d_sel_hyp, d_sel_tok, olen,
                                                           d_r_new, win_start, win_end);
  LAUNCH_CK();
  return LGASR_OK;
}

extern "C" int lgasr_ctc_extend_state(const float* d_xT, int T_old, int T_new, int Tp_old, int Tp_new, int blank, int n,
                                      const float* d_r_old, float* d_r_new, void* stream) {
  if (!d_xT || !d_r_old || !d_r_new || n < 1 || T_old < 1 || T_new < T_old || Tp_old < T_old || Tp_new < T_new)
    return fail(LGASR_EINVAL, "bad argument");
  const size_t smem = (size_t)(T_new - T_old + 1) * sizeof(float);
  if (smem > 48 * 1024) return fail(LGASR_EINVAL, "extension too long");
  k_ctc_extend_state<<<n, 128, smem, (cudaStream_t)stream>>>(d_xT, T_old, T_new, Tp_old, Tp_new, blank, d_r_old, d_r_new);
  LAUNCH_CK();
  return LGASR_OK;
}

// ---- either side of the step: scorer caches by parent, logits -> log-probabilities --------------------------
extern "C" int lgasr_reorder_rows(lgasr_handle* h, void* d_dst, const void* d_src, int64_t n_outer, int64_t dst_outer_bytes,
                                  int64_t src_outer_bytes, int64_t dst_row_bytes, int64_t src_row_bytes, int64_t row_bytes,
                                  void* stream) {
  if (!h) return fail(LGASR_EINVAL, "null handle");
  if (!d_dst || !d_src || n_outer < 1 || n_outer > 65535 || row_bytes < 0 || dst_row_bytes < row_bytes ||
      src_row_bytes < row_bytes || (row_bytes & 3))
    return fail(LGASR_EINVAL, "bad argument");
  if (d_dst == d_src) return fail(LGASR_EINVAL, "in-place reorder is not possible: use two buffers");
  if (row_bytes == 0 || h->step == 0) return LGASR_OK;
  Dev& d = h->d;
  const int BW = d.B * d.W;
  if (BW > 65535) return fail(LGASR_EINVAL, "too many rows");
  const unsigned long long all = (unsigned long long)(uintptr_t)d_dst | (unsigned long long)(uintptr_t)d_src |
                                 (unsigned long long)dst_outer_bytes | (unsigned long long)src_outer_bytes |
                                 (unsigned long long)dst_row_bytes | (unsigned long long)src_row_bytes |
                                 (unsigned long long)row_bytes;
  const bool v16 = (all & 15) == 0;
  const long long nvec = row_bytes / (v16 ? 16 : 4);
  long long gx = (nvec + (long long)kCopyThreads * kCopyUnroll * 2 - 1) / ((long long)kCopyThreads * kCopyUnroll * 2);
  if (gx < 1) gx = 1;
  if (gx > 1024) gx = 1024;
  dim3 grid((unsigned)gx, (unsigned)BW, (unsigned)n_outer);
  cudaStream_t st = (cudaStream_t)stream;
  // the parents of the last step and the running counts it produced (what lgasr_poll would copy out)
  if (v16)
    k_reorder_rows<uint4><<<grid, kCopyThreads, 0, st>>>(d.step_parent, d.n_run[h->pp], d.W, (char*)d_dst, (const char*)d_src,
                                                         dst_outer_bytes, src_outer_bytes, dst_row_bytes, src_row_bytes,
                                                         row_bytes);
  else
    k_reorder_rows<unsigned><<<grid, kCopyThreads, 0, st>>>(d.step_parent, d.n_run[h->pp], d.W, (char*)d_dst,
                                                            (const char*)d_src, dst_outer_bytes, src_outer_bytes,
                                                            dst_row_bytes, src_row_bytes, row_bytes);
  LAUNCH_CK();
  return LGASR_OK;
}

template <bool BF16, bool VECIO>
static int launch_lsm(const void* logits, long long ld_in, float* out, long long ld_out, int rows, int V, int SL, int CS,
                      cudaStream_t st, float* lse_out) {
  auto kern = k_log_softmax_rows<BF16, VECIO>;
  const size_t smem = (size_t)SL * (BF16 ? 2 : 4);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail(LGASR_ECUDA, "%s", cudaGetErrorString(e));
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)CS, (unsigned)rows, 1);
  cfg.blockDim = dim3(kLsmThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)CS;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, logits, ld_in, out, ld_out, V, SL, lse_out);
  if (e != cudaSuccess) return fail(LGASR_ECUDA, "%s", cudaGetErrorString(e));
  g_launches.fetch_add(1);
  return LGASR_OK;
}

static int lsm_impl(const void* d_logits, int dtype, int64_t ld_in, float* d_out, int64_t ld_out, float* d_lse, int rows,
                    int vocab, void* stream) {
  if (!d_logits || (!d_out && !d_lse) || rows < 0 || rows > 65535 || vocab < 1 || ld_in < vocab || (d_out && ld_out < vocab) ||
      (dtype != LGASR_F32 && dtype != LGASR_BF16))
    return fail(LGASR_EINVAL, "bad argument");
  if (rows == 0) return LGASR_OK;
  // slice of a row per CTA: at most 16 K floats (64 KB: three CTAs per SM) -> cluster of 1, 2, 4 or 8 CTAs
  const int kSlice = 16 * 1024;
  int CS = 1;
  while (CS < 8 && (long long)CS * kSlice < vocab) CS <<= 1;
  if ((long long)CS * kSlice < vocab) return fail(LGASR_EINVAL, "vocabulary above %d", 8 * kSlice);
  int SL = (vocab + CS - 1) / CS;
  SL = (SL + 7) & ~7;  // slices start on 32-byte boundaries of the fp32 row (16-byte of a bf16 row)
  const bool bf = dtype == LGASR_BF16;
  const int vq = bf ? 8 : 4;
  const bool vec = (vocab % vq) == 0 && (ld_in % vq) == 0 && (!d_out || (ld_out % 4) == 0) && (((uintptr_t)d_logits) & 15) == 0 &&
                   (!d_out || (((uintptr_t)d_out) & 15) == 0);
  cudaStream_t st = (cudaStream_t)stream;
  if (bf) return vec ? launch_lsm<true, true>(d_logits, ld_in, d_out, ld_out, rows, vocab, SL, CS, st, d_lse)
                     : launch_lsm<true, false>(d_logits, ld_in, d_out, ld_out, rows, vocab, SL, CS, st, d_lse);
  return vec ? launch_lsm<false, true>(d_logits, ld_in, d_out, ld_out, rows, vocab, SL, CS, st, d_lse)
             : launch_lsm<false, false>(d_logits, ld_in, d_out, ld_out, rows, vocab, SL, CS, st, d_lse);
}

extern "C" int lgasr_log_softmax_rows(const void* d_logits, int dtype, int64_t ld_in, float* d_out, int64_t ld_out, int rows,
                                      int vocab, void* stream) {
  if (!d_out) return fail(LGASR_EINVAL, "bad argument");
  return lsm_impl(d_logits, dtype, ld_in, d_out, ld_out, nullptr, rows, vocab, stream);
}

extern "C" int lgasr_row_logsumexp(const void* d_logits, int dtype, int64_t ld_in, float* d_lse, int rows, int vocab,
                                   void* stream) {
  if (!d_lse) return fail(LGASR_EINVAL, "bad argument");
  return lsm_impl(d_logits, dtype, ld_in, nullptr, 0, d_lse, rows, vocab, stream);
}

extern "C" int lgasr_score_filter(const int64_t* d_yseq, int64_t ld_y, int cur_len, int rows, int vocab, int notimestamps,
                                  int first_time, int last_time, int sos, int eos, float* d_out, int64_t ld_out,
                                  void* stream) {
  if (!d_yseq || !d_out || rows < 0 || vocab < 1 || cur_len < 1 || ld_y < cur_len || ld_out < vocab || first_time < 0 ||
      last_time < first_time || last_time >= vocab || eos < 0 || eos >= vocab)
    return fail(LGASR_EINVAL, "bad argument");
  if (rows == 0) return LGASR_OK;
  const int vec_ok = (vocab % 4 == 0) && (ld_out % 4 == 0) && ((((uintptr_t)d_out) & 15) == 0);
  k_score_filter<<<rows, kFilterThreads, 0, (cudaStream_t)stream>>>((const long long*)d_yseq, ld_y, cur_len, vocab,
                                                                     notimestamps, first_time, last_time, sos, eos, d_out,
                                                                     ld_out, vec_ok);
  LAUNCH_CK();
  return LGASR_OK;
}

// ---- lgasr_row_logsumexp_noalloc ---------------------------------------------------------------------------------
// No counterpart in the reference.  A variant of the row-logsumexp kernel
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it) and whose
// launch path cannot use thread-block clusters: a long row is split over plain CTAs and the slices' (max, sum) pairs
// meet in SCRATCH memory instead of the cluster's distributed shared memory.  The scratch is a file-scope device array
// of fixed size: every GPU that loads the library holds it until exit, every call shares it (calls that overlap on the
// GPU race on it), and it keeps the last call's contents - the tickets are cleared at the start of every call.
constexpr int kNoAllocMaxRows = 4096;     // LGASR_NOALLOC_MAX_ROWS
constexpr int kNoAllocMaxSplits = 8;
constexpr int kNoAllocMaxVocab = 1 << 20;  // LGASR_NOALLOC_MAX_VOCAB
constexpr int kNoAllocThreads = 256;
struct NoAllocScratch {
  float2 part[kNoAllocMaxRows][kNoAllocMaxSplits];  // (max, sum of exp relative to it) of every slice: 256 KB
  int ticket[kNoAllocMaxRows];                      // slices of the row that have delivered: 16 KB, zero between calls
};
static __device__ __align__(256) NoAllocScratch g_noalloc_scratch;  // 278 528 bytes per GPU

__device__ __forceinline__ void lse_merge(float& m, float& s, float m2, float s2) {
  // (m, s) (+) (m2, s2) with s = sum exp(x - m); an empty side has m = -inf, s = 0
  const float mm = fmaxf(m, m2);
  if (mm == -INFINITY) return;
  s = s * exp2f((m - mm) * 1.4426950408889634f) + s2 * exp2f((m2 - mm) * 1.4426950408889634f);
  m = mm;
}

// grid (splits, rows).  `scr` is the DEVICE address of g_noalloc_scratch (cudaGetSymbolAddress on the host side).
template <bool BF16, bool VEC>
__global__ void __launch_bounds__(kNoAllocThreads) k_row_logsumexp_scratch(const void* __restrict__ logits, long long ld_in,
                                                                          float* __restrict__ lse_out, int V, int SL,
                                                                          NoAllocScratch* scr) {
  const int r = blockIdx.y, sp = blockIdx.x, nsp = gridDim.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int e0 = sp * SL;
  const int e1 = min(V, e0 + SL);
  float m = -INFINITY, s = 0.f;
  constexpr float kL2E = 1.4426950408889634f;
  // online (max, sum): s = sum of 2^((x - m) * log2 e) relative to the running maximum m (MUFU ex2, as the staged kernel)
  auto take = [&](float x) {
    if (x > m) {
      s = s * ex2_approx((m - x) * kL2E) + 1.f;  // 2^(-inf) = 0 on the first element
      m = x;
    } else if (x > -INFINITY) {
      s += ex2_approx((x - m) * kL2E);
    }
  };
  // four values at once: one maximum update (rare after the first few vectors), then four independent exponentials
  auto take4 = [&](float a, float b, float c, float d4) {
    const float mv = fmaxf(fmaxf(a, b), fmaxf(c, d4));
    if (mv > m) {
      s *= ex2_approx((m - mv) * kL2E);
      m = mv;
    }
    if (m > -INFINITY)
      s += (ex2_approx((a - m) * kL2E) + ex2_approx((b - m) * kL2E)) + (ex2_approx((c - m) * kL2E) + ex2_approx((d4 - m) * kL2E));
  };
  if (BF16) {
    const __nv_bfloat16* row = reinterpret_cast<const __nv_bfloat16*>(logits) + (long long)r * ld_in;
    if (VEC) {  // SL % 8 == 0, 16-byte aligned rows
      const uint4* p = reinterpret_cast<const uint4*>(row + e0);
      const int nq = (e1 - e0) >> 3;
#pragma unroll 2
      for (int q = tid; q < nq; q += kNoAllocThreads) {
        const uint4 u = __ldg(p + q);
        take4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u), __uint_as_float(u.y << 16),
              __uint_as_float(u.y & 0xffff0000u));
        take4(__uint_as_float(u.z << 16), __uint_as_float(u.z & 0xffff0000u), __uint_as_float(u.w << 16),
              __uint_as_float(u.w & 0xffff0000u));
      }
      for (int e = e0 + (nq << 3) + tid; e < e1; e += kNoAllocThreads) take(__bfloat162float(row[e]));
    } else {
      for (int e = e0 + tid; e < e1; e += kNoAllocThreads) take(__bfloat162float(row[e]));
    }
  } else {
    const float* row = reinterpret_cast<const float*>(logits) + (long long)r * ld_in;
    if (VEC) {
      const float4* p = reinterpret_cast<const float4*>(row + e0);
      const int nq = (e1 - e0) >> 2;
#pragma unroll 2
      for (int q = tid; q < nq; q += kNoAllocThreads) {
        const float4 v = __ldg(p + q);
        take4(v.x, v.y, v.z, v.w);
      }
      for (int e = e0 + (nq << 2) + tid; e < e1; e += kNoAllocThreads) take(__ldg(row + e));
    } else {
      for (int e = e0 + tid; e < e1; e += kNoAllocThreads) take(__ldg(row + e));
    }
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
    lse_merge(m, s, m2, s2);
  }
  __shared__ float2 wred[kNoAllocThreads / 32];
  __shared__ int is_last;
  if (lane == 0) wred[warp] = make_float2(m, s);
  __syncthreads();
  if (tid == 0) {
    m = wred[0].x;
    s = wred[0].y;
    for (int w = 1; w < kNoAllocThreads / 32; ++w) lse_merge(m, s, wred[w].x, wred[w].y);
    scr->part[r][sp] = make_float2(m, s);
    __threadfence();  // the pair is visible before the ticket is
    const int t = atomicAdd(&scr->ticket[r], 1);
    is_last = (t == nsp - 1);
    if (is_last) {
      __threadfence();
      // the last slice to arrive combines all of them, in slice order (the result does not depend on arrival order)
      float M = -INFINITY, S = 0.f;
      for (int i = 0; i < nsp; ++i) {
        const float2 p = __ldcg(&scr->part[r][i]);
        lse_merge(M, S, p.x, p.y);
      }
      lse_out[r] = (M == -INFINITY) ? -INFINITY : M + logf(S);
      scr->ticket[r] = 0;  // left clean for the next call (which clears the tickets again anyway)
    }
  }
}

// Device address of g_noalloc_scratch, resolved once per device with that device current.  (In host code the array's
// name is only a stand-in for the symbol: handing it to a kernel would fault or touch the wrong memory.)
static NoAllocScratch* noalloc_scratch_ptr(int* rc) {
  static std::mutex mu;
  static NoAllocScratch* ptr[64] = {};
  int dev = 0;
  *rc = LGASR_OK;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    *rc = fail(LGASR_ECUDA, "cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return nullptr;
  }
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      *rc = fail(LGASR_ECUDA, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    ptr[dev] = static_cast<NoAllocScratch*>(p);
  }
  return ptr[dev];
}

extern "C" int lgasr_row_logsumexp_noalloc(const void* d_logits, int dtype, int64_t ld_in, float* d_lse, int rows, int vocab,
                                           void* stream) {
  if (!d_logits || !d_lse || rows < 0 || vocab < 1 || ld_in < vocab || (dtype != LGASR_F32 && dtype != LGASR_BF16))
    return fail(LGASR_EINVAL, "bad argument");
  if (rows > kNoAllocMaxRows)
    return fail(LGASR_EINVAL, "lgasr_row_logsumexp_noalloc: %d rows exceed the fixed scratch (LGASR_NOALLOC_MAX_ROWS = %d)", rows,
                kNoAllocMaxRows);
  if (vocab > kNoAllocMaxVocab)
    return fail(LGASR_EINVAL, "lgasr_row_logsumexp_noalloc: vocabulary %d exceeds LGASR_NOALLOC_MAX_VOCAB = %d", vocab,
                kNoAllocMaxVocab);
  if (rows == 0) return LGASR_OK;
  int rc = LGASR_OK;
  NoAllocScratch* scr = noalloc_scratch_ptr(&rc);
  if (!scr) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // slices of at least 8 K entries, at most kNoAllocMaxSplits of them, starting on 32-byte boundaries
  int nsp = (vocab + 8191) / 8192;
  if (nsp > kNoAllocMaxSplits) nsp = kNoAllocMaxSplits;
  if (nsp < 1) nsp = 1;
  int SL = (vocab + nsp - 1) / nsp;
  SL = (SL + 7) & ~7;
  nsp = (vocab + SL - 1) / SL;
  const bool bf = dtype == LGASR_BF16;
  const int vq = bf ? 8 : 4;
  const bool vec = (ld_in % vq) == 0 && ((((uintptr_t)d_logits) & 15) == 0);
  // the tickets must start at zero whatever an earlier (possibly failed) call left behind; no allocation involved
  CK(cudaMemsetAsync(&scr->ticket[0], 0, sizeof(int) * (size_t)rows, st));
  const dim3 grid(nsp, rows);
  if (bf) {
    if (vec) k_row_logsumexp_scratch<true, true><<<grid, kNoAllocThreads, 0, st>>>(d_logits, ld_in, d_lse, vocab, SL, scr);
    else k_row_logsumexp_scratch<true, false><<<grid, kNoAllocThreads, 0, st>>>(d_logits, ld_in, d_lse, vocab, SL, scr);
  } else {
    if (vec) k_row_logsumexp_scratch<false, true><<<grid, kNoAllocThreads, 0, st>>>(d_logits, ld_in, d_lse, vocab, SL, scr);
    else k_row_logsumexp_scratch<false, false><<<grid, kNoAllocThreads, 0, st>>>(d_logits, ld_in, d_lse, vocab, SL, scr);
  }
  LAUNCH_CK();
  return LGASR_OK;
}
