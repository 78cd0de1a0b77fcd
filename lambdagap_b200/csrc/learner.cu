// learner.cu -- host side of the device tree learner + its C ABI.
//
// Mirrors SerialTreeLearner::Train (src/treelearner/serial_tree_learner.cpp:179-245) and, with a communicator,
// DataParallelTreeLearner (src/treelearner/data_parallel_tree_learner.cpp:123-463).  The leaf-wise loop is run in
// WAVES: all leaves that may still be needed are split together (partition -> histograms of the smaller children ->
// scan of all children -> selection of the next wave), and the reference's one-split-at-a-time order is replayed at
// the end.  Every kernel finds its work in device memory; the whole tree is one CUDA graph with a
// WHILE conditional node around the wave (GBL_USE_GRAPH=0, stage profiling and the data-parallel learner drive the
// waves from the host, reading the small tree state back once per wave).  The reference CUDA learner synchronises with the host
// twice per split (src/treelearner/cuda/cuda_single_gpu_tree_learner.cpp:246-325).
#include <unistd.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "comm.cuh"
#include "kernels.cuh"

namespace gbl {

struct Learner {
  const Dataset* ds = nullptr;
  gbl_learner_config cfg{};
  Comm* comm = nullptr;
  cudaStream_t stream = nullptr;
  LearnerDev dev{};
  // owned device memory
  TreeState* ts = nullptr;
  Node* nodes = nullptr;
  unsigned long long* bkey = nullptr;
  unsigned char* expanded = nullptr;
  Expand* wave = nullptr;
  unsigned* job_done = nullptr;
  int* unit_left = nullptr;
  unsigned* unit_bits = nullptr;
  int* leaf_node = nullptr;
  int* split_node = nullptr;
  int* seg_node = nullptr;
  int* seg_unit_begin = nullptr;
  double* renew_sums = nullptr;             // [2 * num_leaves] true gradient sums per leaf (quant_train_renew_leaf)
  unsigned char* replay_scratch = nullptr;  // only for trees too large for the replay's shared memory
  int* idx[3] = {nullptr, nullptr, nullptr};
  uint8_t* rec = nullptr;  // row records (bins + gradient pair), rebuilt when the gradient format changes
  bool rec_quant = false;
  bool use_mc = false;            // gbl_learner_set_monotone
  double monotone_penalty = 0.0;
  int max_cat_threshold = 32, max_cat_to_onehot = 4, min_data_per_group = 100;  // gbl_learner_set_categorical
  double cat_l2 = 10.0, cat_smooth = 10.0;
  int* cat_features = nullptr;  // device list of the categorical features
  int num_cat_features = 0;
  short* gh_q = nullptr;
  unsigned short* leaf_u16 = nullptr;  // leaf id per row, scratch of the score update
  long long* hist_pool = nullptr;
  size_t hist_pool_words = 0;
  unsigned char* splittable = nullptr;
  signed char* feature_mask = nullptr;
  BestSplit* cand = nullptr;
  gbl_split_record* records = nullptr;
  double* leaf_values = nullptr;
  double* leaf_values_tmp = nullptr;
  double* rand_g = nullptr;
  double* rand_h = nullptr;
  int64_t rand_n = 0;
  int rand_start = 0;
  int* bag_indices = nullptr;
  int bag_count = 0;
  int* oob_indices = nullptr;  // rows outside the bag (complement of bag_indices)
  int oob_count = 0;
  // peer-memory exchange of the data-parallel learner (exchange.cu): control block, device pointer tables, mapped peers
  bool peer = false;
  unsigned char* xchg_ctl = nullptr;     // own: flags[R] | maxbuf[R][2] | epoch | dead
  void** xchg_tables = nullptr;          // device: pool[R] | flags[R] | maxbuf[R]
  std::vector<void*> peer_pool_maps;     // cudaIpcOpenMemHandle results (other ranks only)
  std::vector<void*> peer_ctl_maps;
  // device-drawn bags (gbl_learner_bagging_dev): per-block generator states, 1 bit per row, block counts / offsets,
  // {bag size, out-of-bag size} on the device and in pinned host memory
  bool bag_on_device = false;
  unsigned* bag_state = nullptr;
  unsigned* bag_bits = nullptr;
  int* bag_blk_cnt = nullptr;
  int* bag_blk_off = nullptr;
  int* bag_counts_dev = nullptr;
  int* bag_counts_pinned = nullptr;
  int bag_seed = 0;
  float* bag_label = nullptr;  // balanced bagging: the labels that pick a row's fraction
  int* bag_row_query = nullptr;     // bagging by query: query of every row
  unsigned* bag_qbits = nullptr;    // 1 bit per query
  int* bag_qblk_cnt = nullptr;
  int bag_num_queries = 0;
  float* grad_stage = nullptr;  // train_host staging
  float* hess_stage = nullptr;
  long long* tap_hist = nullptr;
  // pinned host mirrors
  gbl_split_record* records_pinned = nullptr;
  TreeState* ts_pinned = nullptr;
  // graph
  bool use_graph = true;
  cudaGraphExec_t graph_exec = nullptr;
  cudaGraphConditionalHandle graph_cond = 0;
  const float* graph_grad = nullptr;
  const float* graph_hess = nullptr;
  int graph_rand_start = -1;
  const int* graph_bag = nullptr;
  int graph_bag_count = -1;
  bool tree_pending = false;
  bool has_tree = false;
  // profiling (CUDA events around every stage)
  bool profiling = false;
  std::vector<cudaEvent_t> ev;     // 2 per timed stage instance
  std::vector<int> ev_stage;       // stage id per event pair
  size_t ev_used = 0;
  gbl_profile prof{};
  cudaEvent_t timer[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t state_ready = nullptr;  // host-driven loop: the tree state of the last selection has reached the host
  bool tree_in_graph = false;        // the pending tree runs as a graph replay (launches counted at train_finish)
};

enum Stage { kStageHist = 0, kStageFind = 1, kStagePart = 2, kStageOther = 3, kStageComm = 4 };

struct StageTimer {
  Learner* L;
  cudaStream_t s;
  size_t slot = 0;
  bool on;
  StageTimer(Learner* L_, cudaStream_t s_, int stage) : L(L_), s(s_), on(L_->profiling) {
    if (!on) return;
    if (L->ev_used + 2 > L->ev.size()) {
      for (int k = 0; k < 2; ++k) {
        cudaEvent_t e;
        GBL_CUDA(cudaEventCreate(&e));
        L->ev.push_back(e);
      }
      L->ev_stage.push_back(stage);
    } else {
      L->ev_stage[L->ev_used / 2] = stage;
    }
    slot = L->ev_used;
    L->ev_used += 2;
    record(L->ev[slot]);
  }
  ~StageTimer() {
    if (on) record(L->ev[slot + 1]);
  }
  // Inside a stream capture the record must be an *external* event node, otherwise the event belongs to the
  // capture and cannot be queried after the graph has run.
  void record(cudaEvent_t e) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(s, &st);
    cudaEventRecordWithFlags(e, s, st == cudaStreamCaptureStatusActive ? cudaEventRecordExternal : cudaEventRecordDefault);
  }
};

static void validate_config(const gbl_learner_config& c) {
  if (c.num_leaves < 2) fatal("learner: num_leaves must be >= 2 (got %d)", c.num_leaves);
  if (c.num_leaves > 32768) fatal("learner: num_leaves=%d too large", c.num_leaves);
  if (c.min_data_in_leaf < 0) fatal("learner: min_data_in_leaf must be >= 0");
  if (c.use_quantized_grad && (c.num_grad_quant_bins < 2 || c.num_grad_quant_bins > 126))
    fatal("learner: num_grad_quant_bins=%d unsupported (2..126)", c.num_grad_quant_bins);
  if (c.lambda_l1 < 0 || c.lambda_l2 < 0) fatal("learner: lambda_l1 / lambda_l2 must be >= 0");
}

static LearnerParams make_params(const Learner* L) {
  LearnerParams p{};
  const gbl_learner_config& c = L->cfg;
  p.num_leaves = c.num_leaves;
  p.max_depth = c.max_depth;
  p.min_data_in_leaf = c.min_data_in_leaf;
  p.quantized = c.use_quantized_grad ? 1 : 0;
  p.num_grad_quant_bins = c.num_grad_quant_bins;
  p.is_constant_hessian = c.is_constant_hessian ? 1 : 0;
  p.num_features = L->ds->num_features;
  p.num_tiles = L->ds->num_tiles;
  p.num_ranks = L->comm != nullptr ? L->comm->num_ranks : 1;
  p.use_mc = L->use_mc ? 1 : 0;
  p.monotone_penalty = L->monotone_penalty;
  p.max_cat_threshold = L->max_cat_threshold;
  p.max_cat_to_onehot = L->max_cat_to_onehot;
  p.min_data_per_group = L->min_data_per_group;
  p.cat_l2 = L->cat_l2;
  p.cat_smooth = L->cat_smooth;
  p.num_cat_features = L->num_cat_features;
  p.min_sum_hessian_in_leaf = c.min_sum_hessian_in_leaf;
  p.lambda_l1 = c.lambda_l1;
  p.lambda_l2 = c.lambda_l2;
  p.min_gain_to_split = c.min_gain_to_split;
  p.max_delta_step = c.max_delta_step;
  p.path_smooth = c.path_smooth;
  return p;
}

static void setup_peer(Learner* L);
static void teardown_peer(Learner* L);
static void apply_gh_window_to_stream(Learner* L);

static void drop_graph(Learner* L) {
  if (L->graph_exec != nullptr) {
    cudaGraphExecDestroy(L->graph_exec);
    L->graph_exec = nullptr;
  }
}

static void free_tree_buffers(Learner* L) {
  void* ptrs[] = {L->nodes, L->bkey, L->expanded, L->wave, L->job_done, L->leaf_node, L->split_node, L->seg_node,
                  L->seg_unit_begin, L->unit_left, L->unit_bits, L->hist_pool, L->splittable, L->cand, L->records,
                  L->leaf_values, L->leaf_values_tmp, L->replay_scratch, L->renew_sums};
  for (void* p : ptrs) cudaFree(p);
  if (L->records_pinned) cudaFreeHost(L->records_pinned);
  L->nodes = nullptr;
  L->bkey = nullptr;
  L->expanded = nullptr;
  L->wave = nullptr;
  L->job_done = nullptr;
  L->leaf_node = nullptr;
  L->split_node = nullptr;
  L->seg_node = nullptr;
  L->seg_unit_begin = nullptr;
  L->unit_left = nullptr;
  L->unit_bits = nullptr;
  L->hist_pool = nullptr;
  L->splittable = nullptr;
  L->cand = nullptr;
  L->records = nullptr;
  L->leaf_values = nullptr;
  L->leaf_values_tmp = nullptr;
  L->replay_scratch = nullptr;
  L->renew_sums = nullptr;
  L->records_pinned = nullptr;
}

// Row records.  The histogram kernel gathers, per row and 32-feature tile, the tile's bins and the row's gradient
// pair.  Float mode: DRAM serves whole bursts, so two separate gathers (a 32-byte bin sector here, an 8-byte pair
// there, out of an 8 x num_data byte array that does not stay in L2) cost two bursts -- measured 181 B of DRAM traffic
// per row against 40 B algorithmic on the Higgs shape.  A float-mode record keeps them together, one block per tile:
//   [32 bytes of bins of tile 0][int2 fixed-point pair + padding] [32 bytes of bins of tile 1][pair + padding] ...
// (64 B per tile; every tile block has its own copy of the pair because every tile is accumulated by its own CTAs).
// The records belong to the learner, not the dataset: gradients are per learner, several learners may share a dataset.
// Quantized mode: the int8 pairs are 2 bytes per row -- an array that stays in L2 (21 MB on the Higgs shape) -- so they
// live in gh_q and the "records" are the dataset's read-only row mirror itself (32 bytes per row and tile, no copy).
// With a single tile whose unused tail has room for the pair (at most 24 features) the record is just the 32-byte sector.
static void record_layout(int num_features, int num_tiles, int* gh_off, int* tile_stride, int* stride) {
  const int inside = static_cast<int>(round_up(num_features, 8));
  if (num_tiles == 1 && inside + 8 <= kTile) {
    *gh_off = inside;
    *tile_stride = kTile;
  } else {
    *gh_off = kTile;
    *tile_stride = 2 * kTile;
  }
  *stride = *tile_stride * num_tiles;
}

static void build_records(Learner* L) {
  const Dataset* ds = L->ds;
  const bool quant = L->cfg.use_quantized_grad != 0;
  if (quant) {
    cudaFree(L->rec);
    L->rec = nullptr;
    L->rec_quant = true;
    L->dev.rec = ds->row_bins;  // never written in this mode
    L->dev.rec_stride = ds->row_stride;
    L->dev.tile_stride = kTile;
    L->dev.gh_off = 0;
    apply_gh_window_to_stream(L);
    return;
  }
  int gh_off = 0, tile_stride = 0, stride = 0;
  record_layout(ds->num_features, ds->num_tiles, &gh_off, &tile_stride, &stride);
  if (L->rec != nullptr && !L->rec_quant) return;
  cudaFree(L->rec);
  L->rec = nullptr;
  GBL_CUDA(cudaMalloc(&L->rec, static_cast<size_t>(ds->num_data) * stride + 64));
  GBL_CUDA(cudaMemsetAsync(L->rec, 0, static_cast<size_t>(ds->num_data) * stride + 64, L->stream));
  launch_build_records(L->rec, stride, tile_stride, ds->row_bins, ds->row_stride, static_cast<int>(ds->num_data), L->stream);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  L->rec_quant = false;
  L->dev.rec = L->rec;
  L->dev.rec_stride = stride;
  L->dev.tile_stride = tile_stride;
  L->dev.gh_off = gh_off;
  apply_gh_window_to_stream(L);
}

// (re)allocates everything that depends on the config.
//
// Capacities.  K = num_leaves - 1 splits make the final tree.  A wave may expand up to K nodes; waves expand
// speculatively until spec_limit nodes have been expanded, afterwards one certain node per wave (at most K more).
// Hence max_expand = spec_limit + 2K - 1 expansions, one fresh histogram slot per expansion, two nodes per expansion.
// spec_limit is 4K unless the histogram pool would not fit the memory budget (GBL_HIST_POOL_MB, default a quarter
// of the device's free memory), in which case speculation is cut back -- down to none (the reference's own order).
static void alloc_tree_buffers(Learner* L) {
  const Dataset* ds = L->ds;
  const int nl = L->cfg.num_leaves;
  const int K = nl - 1;
  const long long entry_words = L->cfg.use_quantized_grad ? 1 : 2;
  const long long slot_stride = static_cast<long long>(ds->num_tiles) * kMaxBin * kTile * entry_words + kTailWords;
  size_t free_b = 0, total_b = 0;
  GBL_CUDA(cudaMemGetInfo(&free_b, &total_b));
  long long budget = static_cast<long long>(free_b / 4);
  if (const char* e = std::getenv("GBL_HIST_POOL_MB")) budget = std::max(1LL, std::atoll(e)) << 20;
  const long long slot_bytes = slot_stride * 8;
  long long spec = 4LL * K;
  if (const char* e = std::getenv("GBL_SPEC_LIMIT")) spec = std::max(0LL, std::min(spec, std::atoll(e)));
  while (spec > 0 && (spec + 2LL * K) * slot_bytes > budget) spec /= 2;
  const int max_expand = static_cast<int>(spec > 0 ? spec + 2LL * K - 1 : K);
  const int max_slots = max_expand + 1;
  const int max_nodes = 2 * max_expand + 1;
  const int max_cand = K;
  const int max_part_units = static_cast<int>(ds->num_data / kPartUnit) + max_cand + 2;
  L->hist_pool_words = static_cast<size_t>(slot_stride) * max_slots;
  GBL_CUDA(cudaMalloc(&L->nodes, sizeof(Node) * max_nodes));
  GBL_CUDA(cudaMalloc(&L->bkey, sizeof(unsigned long long) * max_nodes));
  GBL_CUDA(cudaMalloc(&L->expanded, max_nodes));
  GBL_CUDA(cudaMalloc(&L->wave, sizeof(Expand) * (max_cand + 1)));
  GBL_CUDA(cudaMalloc(&L->job_done, sizeof(unsigned) * 2 * max_cand));
  GBL_CUDA(cudaMalloc(&L->leaf_node, sizeof(int) * nl));
  GBL_CUDA(cudaMalloc(&L->split_node, sizeof(int) * nl));
  GBL_CUDA(cudaMalloc(&L->seg_node, sizeof(int) * max_slots));
  GBL_CUDA(cudaMalloc(&L->seg_unit_begin, sizeof(int) * (max_slots + 1)));
  GBL_CUDA(cudaMalloc(&L->unit_left, sizeof(int) * (max_part_units + 1)));
  GBL_CUDA(cudaMalloc(&L->unit_bits, sizeof(unsigned) * static_cast<size_t>(max_part_units) * (kPartUnit / 32)));
  GBL_CUDA(cudaMalloc(&L->hist_pool, sizeof(long long) * L->hist_pool_words));
  GBL_CUDA(cudaMalloc(&L->splittable, static_cast<size_t>(max_slots) * ds->num_features));
  GBL_CUDA(cudaMalloc(&L->cand, sizeof(BestSplit) * 2 * static_cast<size_t>(max_cand) * ds->num_features));
  GBL_CUDA(cudaMalloc(&L->records, sizeof(gbl_split_record) * nl));
  GBL_CUDA(cudaMemset(L->records, 0, sizeof(gbl_split_record) * nl));
  GBL_CUDA(cudaMalloc(&L->leaf_values, sizeof(double) * nl));
  GBL_CUDA(cudaMalloc(&L->leaf_values_tmp, sizeof(double) * nl));
  GBL_CUDA(cudaMalloc(&L->renew_sums, sizeof(double) * 2 * nl));
  GBL_CUDA(cudaMallocHost(&L->records_pinned, sizeof(gbl_split_record) * nl));
  LearnerDev& d = L->dev;
  d.nodes = L->nodes;
  d.bkey = L->bkey;
  d.expanded = L->expanded;
  d.wave = L->wave;
  d.job_done = L->job_done;
  d.leaf_node = L->leaf_node;
  d.split_node = L->split_node;
  d.seg_node = L->seg_node;
  d.seg_unit_begin = L->seg_unit_begin;
  d.unit_left = L->unit_left;
  d.unit_bits = L->unit_bits;
  d.hist_pool = L->hist_pool;
  d.slot_stride = slot_stride;
  d.splittable = L->splittable;
  d.cand = L->cand;
  d.records = L->records;
  d.leaf_values = L->leaf_values;
  d.max_nodes = max_nodes;
  d.max_slots = max_slots;
  d.max_cand = max_cand;
  d.max_expand = max_expand;
  d.spec_limit = static_cast<int>(spec);
  d.max_part_units = max_part_units;
  // measured on the Higgs shape over 240 boosting iterations (early trees: balanced; late trees: flat gains, long
  // chains): 0.5 / 32 is the flat optimum between wasted expansions (larger window) and extra waves (smaller window)
  d.spec_rho = 0.5f;
  d.spec_min = 32;
  if (const char* e = std::getenv("GBL_SPEC_RHO")) d.spec_rho = std::min(1.0f, std::max(0.05f, static_cast<float>(std::atof(e))));
  if (const char* e = std::getenv("GBL_SPEC_MIN")) d.spec_min = std::max(1, std::atoi(e));
  d.p = make_params(L);
  if (finalize_scratch_bytes(d) > kFinalizeSmemMax) GBL_CUDA(cudaMalloc(&L->replay_scratch, finalize_scratch_bytes(d)));
  d.replay_scratch = L->replay_scratch;
}

static Learner* learner_create(const Dataset* ds, const gbl_learner_config* cfg, Comm* comm) {
  if (ds == nullptr || cfg == nullptr) fatal("gbl_learner_create: null argument");
  validate_config(*cfg);
  if (comm != nullptr && comm->device != ds->device) fatal("learner: communicator and dataset on different devices");
  DeviceGuard guard(ds->device);
  cudaDeviceProp prop;
  GBL_CUDA(cudaGetDeviceProperties(&prop, ds->device));
  if (prop.major != 10) fatal("learner: device %d is sm_%d%d; this library is built for sm_100a (B200) only", ds->device,
                              prop.major, prop.minor);
  Learner* L = new Learner();
  L->ds = ds;
  L->cfg = *cfg;
  L->comm = comm;
  const char* g = std::getenv("GBL_USE_GRAPH");
  L->use_graph = !(g != nullptr && g[0] == '0');
  GBL_CUDA(cudaStreamCreateWithFlags(&L->stream, cudaStreamNonBlocking));
  const size_t n = static_cast<size_t>(ds->num_data);
  GBL_CUDA(cudaMalloc(&L->ts, sizeof(TreeState)));
  GBL_CUDA(cudaMemset(L->ts, 0, sizeof(TreeState)));
  GBL_CUDA(cudaMalloc(&L->idx[0], sizeof(int) * (n + 32)));
  GBL_CUDA(cudaMalloc(&L->idx[1], sizeof(int) * (n + 32)));
  GBL_CUDA(cudaMalloc(&L->idx[2], sizeof(int) * (n + 32)));
  launch_iota(L->idx[2], static_cast<int>(n), L->stream);
  GBL_CUDA(cudaMalloc(&L->gh_q, sizeof(short) * n));
  GBL_CUDA(cudaMalloc(&L->leaf_u16, sizeof(unsigned short) * (n + 8)));
  GBL_CUDA(cudaMallocHost(&L->ts_pinned, sizeof(TreeState)));
  LearnerDev& d = L->dev;
  d.col_bins = ds->col_bins;
  d.num_data = static_cast<int>(ds->num_data);
  d.meta = ds->meta_dev;
  {
    std::vector<int> cats;
    for (int f = 0; f < ds->num_features; ++f) {
      if (ds->meta_host[f].is_categorical) cats.push_back(f);
    }
    L->num_cat_features = static_cast<int>(cats.size());
    if (!cats.empty()) {
      GBL_CUDA(cudaMalloc(&L->cat_features, sizeof(int) * cats.size()));
      GBL_CUDA(cudaMemcpy(L->cat_features, cats.data(), sizeof(int) * cats.size(), cudaMemcpyHostToDevice));
    }
    d.cat_features = L->cat_features;
  }
  d.ts = L->ts;
  d.idx[0] = L->idx[0];
  d.idx[1] = L->idx[1];
  d.idx[2] = L->idx[2];
  d.gh_q = L->gh_q;
  d.feature_mask = nullptr;
  d.rand_g = nullptr;
  d.rand_h = nullptr;
  d.rand_start = 0;
  alloc_tree_buffers(L);
  build_records(L);
  histogram_setup();
  setup_peer(L);
  return L;
}

static void learner_free(Learner* L) {
  if (L == nullptr) return;
  DeviceGuard guard(L->ds->device);
  cudaStreamSynchronize(L->stream);
  drop_graph(L);
  teardown_peer(L);
  free_tree_buffers(L);
  cudaFree(L->ts);
  cudaFree(L->idx[0]);
  cudaFree(L->idx[1]);
  cudaFree(L->idx[2]);
  cudaFree(L->rec);
  cudaFree(L->gh_q);
  cudaFree(L->leaf_u16);
  cudaFree(L->feature_mask);
  cudaFree(L->cat_features);
  cudaFree(L->rand_g);
  cudaFree(L->rand_h);
  cudaFree(L->bag_indices);
  cudaFree(L->oob_indices);
  cudaFree(L->bag_state);
  cudaFree(L->bag_bits);
  cudaFree(L->bag_blk_cnt);
  cudaFree(L->bag_blk_off);
  cudaFree(L->bag_counts_dev);
  cudaFree(L->bag_label);
  cudaFree(L->bag_row_query);
  cudaFree(L->bag_qbits);
  cudaFree(L->bag_qblk_cnt);
  if (L->bag_counts_pinned) cudaFreeHost(L->bag_counts_pinned);
  cudaFree(L->grad_stage);
  cudaFree(L->hess_stage);
  cudaFree(L->tap_hist);
  if (L->ts_pinned) cudaFreeHost(L->ts_pinned);
  for (cudaEvent_t e : L->ev) cudaEventDestroy(e);
  if (L->state_ready != nullptr) cudaEventDestroy(L->state_ready);
  for (cudaEvent_t e : L->timer) {
    if (e != nullptr) cudaEventDestroy(e);
  }
  cudaStreamDestroy(L->stream);
  delete L;
}

// ---- peer-memory exchange set-up (data-parallel learner, one process per GPU).  Collective: every rank calls it at the
// same point (learner creation / reset_config).  Handles travel through one int64 all-reduce (each rank fills its own
// 128 bytes of a zeroed table).  If any rank cannot map a peer (ranks sharing one process, no peer access) all ranks
// keep the NCCL path.
constexpr size_t kCtlFlagsOff = 0, kCtlMaxOff = 256, kCtlEpochOff = 1024, kCtlDeadOff = 1028, kCtlDoneOff = 1032, kCtlGoOff = 1036, kCtlBytes = 2048,
                 kCtlAlloc = 2u << 20;

static void teardown_peer(Learner* L) {
  for (void* p : L->peer_pool_maps) cudaIpcCloseMemHandle(p);
  for (void* p : L->peer_ctl_maps) cudaIpcCloseMemHandle(p);
  L->peer_pool_maps.clear();
  L->peer_ctl_maps.clear();
  cudaFree(L->xchg_ctl);
  cudaFree(L->xchg_tables);
  L->xchg_ctl = nullptr;
  L->xchg_tables = nullptr;
  L->peer = false;
  L->dev.px = PeerX{};
}

static void setup_peer(Learner* L) {
  teardown_peer(L);
  Comm* c = L->comm;
  if (c == nullptr || c->num_ranks <= 1) return;
  const int R = c->num_ranks, me = c->rank;
  bool want = R <= 16 && !(L->cfg.use_quantized_grad && L->cfg.quant_train_renew_leaf);
  if (const char* e = std::getenv("GBL_PEER_EXCHANGE")) want = want && std::atoi(e) != 0;
  cudaStream_t s = L->stream;
  // 2 MiB: an allocation of its own (small cudaMalloc blocks share a 2 MiB region, and one exported region can be
  // mapped only once per peer process -- a second learner's control block in the same region could not be opened)
  GBL_CUDA(cudaMalloc(&L->xchg_ctl, kCtlAlloc));
  GBL_CUDA(cudaMemset(L->xchg_ctl, 0, kCtlBytes));
  // kWords int64 words per rank: two 64-byte IPC handles (pool, control block), then process id, the two raw device
  // pointers and the device ordinal -- ranks that share this process (one Booster driving several GPUs from host
  // threads) cannot open each other's IPC handles; they use the raw pointers after enabling peer access instead
  constexpr int kWords = 20;
  long long* table_dev = nullptr;
  GBL_CUDA(cudaMalloc(&table_dev, sizeof(long long) * kWords * R));
  GBL_CUDA(cudaMemset(table_dev, 0, sizeof(long long) * kWords * R));
  GBL_CUDA(cudaDeviceSynchronize());  // the control block is zero before any peer can learn its handle
  cudaIpcMemHandle_t hp, hc;
  bool ok = want;
  if (ok && cudaIpcGetMemHandle(&hp, L->hist_pool) != cudaSuccess) ok = false;
  if (ok && cudaIpcGetMemHandle(&hc, L->xchg_ctl) != cudaSuccess) ok = false;
  cudaGetLastError();
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (ok) {
    long long mine[kWords];
    std::memcpy(mine, &hp, 64);
    std::memcpy(mine + 8, &hc, 64);
    mine[16] = static_cast<long long>(getpid());
    mine[17] = static_cast<long long>(reinterpret_cast<uintptr_t>(L->hist_pool));
    mine[18] = static_cast<long long>(reinterpret_cast<uintptr_t>(L->xchg_ctl));
    mine[19] = L->ds->device;
    GBL_CUDA(cudaMemcpy(table_dev + kWords * me, mine, sizeof(mine), cudaMemcpyHostToDevice));
  }
  comm_allreduce_i64(c, table_dev, static_cast<size_t>(kWords) * R, RedOp::kSum, s);
  GBL_CUDA(cudaStreamSynchronize(s));
  std::vector<long long> table(static_cast<size_t>(kWords) * R);
  GBL_CUDA(cudaMemcpy(table.data(), table_dev, sizeof(long long) * table.size(), cudaMemcpyDeviceToHost));
  cudaFree(table_dev);
  std::vector<void*> pools(R, nullptr), ctls(R, nullptr);
  pools[me] = L->hist_pool;
  ctls[me] = L->xchg_ctl;
  for (int r = 0; r < R && ok; ++r) {
    if (r == me) continue;
    cudaIpcMemHandle_t a, b;
    std::memcpy(&a, &table[kWords * r], 64);
    std::memcpy(&b, &table[kWords * r + 8], 64);
    bool zero = true;
    for (int k = 0; k < kWords; ++k) zero = zero && table[kWords * r + k] == 0;
    if (zero) {  // that rank could not export its memory
      ok = false;
      break;
    }
    if (table[kWords * r + 16] == static_cast<long long>(getpid())) {
      // same process: the peer's allocations are ordinary pointers of this address space
      const int peer_dev = static_cast<int>(table[kWords * r + 19]);
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, L->ds->device, peer_dev) != cudaSuccess || !can) {
        ok = false;
        break;
      }
      const cudaError_t e = cudaDeviceEnablePeerAccess(peer_dev, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        ok = false;
        break;
      }
      cudaGetLastError();
      pools[r] = reinterpret_cast<void*>(static_cast<uintptr_t>(table[kWords * r + 17]));
      ctls[r] = reinterpret_cast<void*>(static_cast<uintptr_t>(table[kWords * r + 18]));
      continue;
    }
    void* pp = nullptr;
    void* pc = nullptr;
    if (cudaIpcOpenMemHandle(&pp, a, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      ok = false;
      break;
    }
    L->peer_pool_maps.push_back(pp);
    if (cudaIpcOpenMemHandle(&pc, b, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      ok = false;
      break;
    }
    L->peer_ctl_maps.push_back(pc);
    pools[r] = pp;
    ctls[r] = pc;
  }
  cudaGetLastError();
  // unanimous or not at all
  float* flag_dev = nullptr;
  GBL_CUDA(cudaMalloc(&flag_dev, sizeof(float)));
  const float mine = ok ? 0.f : 1.f;
  GBL_CUDA(cudaMemcpy(flag_dev, &mine, sizeof(float), cudaMemcpyHostToDevice));
  comm_allreduce_f32(c, flag_dev, 1, RedOp::kMax, s);
  GBL_CUDA(cudaStreamSynchronize(s));
  float any_failed = 1.f;
  GBL_CUDA(cudaMemcpy(&any_failed, flag_dev, sizeof(float), cudaMemcpyDeviceToHost));
  cudaFree(flag_dev);
  if (std::getenv("GBL_DEBUG_PEER") != nullptr)
    std::fprintf(stderr, "[gbl] rank %d/%d: peer-memory exchange %s (wanted %d, mapped here %d)\n", me, R,
                 any_failed != 0.f ? "OFF -> NCCL" : "on", want ? 1 : 0, ok ? 1 : 0);
  if (any_failed != 0.f) {
    teardown_peer(L);
    return;
  }
  std::vector<void*> tables(static_cast<size_t>(3) * R);
  for (int r = 0; r < R; ++r) {
    tables[r] = pools[r];
    tables[R + r] = static_cast<unsigned char*>(ctls[r]) + kCtlFlagsOff;
    tables[2 * R + r] = static_cast<unsigned char*>(ctls[r]) + kCtlMaxOff;
  }
  GBL_CUDA(cudaMalloc(&L->xchg_tables, sizeof(void*) * tables.size()));
  GBL_CUDA(cudaMemcpy(L->xchg_tables, tables.data(), sizeof(void*) * tables.size(), cudaMemcpyHostToDevice));
  PeerX& x = L->dev.px;
  x.pool = reinterpret_cast<long long* const*>(L->xchg_tables);
  x.flags = reinterpret_cast<unsigned* const*>(L->xchg_tables + R);
  x.maxbuf = reinterpret_cast<float* const*>(L->xchg_tables + 2 * R);
  x.epoch = reinterpret_cast<unsigned*>(L->xchg_ctl + kCtlEpochOff);
  x.dead = reinterpret_cast<unsigned*>(L->xchg_ctl + kCtlDeadOff);
  x.done = reinterpret_cast<unsigned*>(L->xchg_ctl + kCtlDoneOff);
  x.go = reinterpret_cast<unsigned*>(L->xchg_ctl + kCtlGoOff);
  x.rank = me;
  double timeout_s = 60.0;
  if (const char* e = std::getenv("GBL_PEER_TIMEOUT_S")) timeout_s = std::max(0.001, std::atof(e));
  int khz = 1900000;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, L->ds->device);
  x.timeout_cycles = static_cast<long long>(timeout_s * 1e3 * khz);
  L->peer = true;
}

// Quantized mode: the int8 gradient pairs (2 bytes per row) are gathered by row id in every wave.  The row records
// stream through L2 at several hundred MB per launch and would evict them (measured: +13 B of DRAM traffic per
// gathered row), so the array is pinned with a persisting access-policy window -- on the learner's stream for the
// host-driven loop and on every kernel node of the tree graph.
static bool gh_window(const Learner* L, cudaAccessPolicyWindow* w) {
  if (!L->cfg.use_quantized_grad || L->gh_q == nullptr) return false;
  if (const char* e = std::getenv("GBL_L2_PERSIST")) {
    if (e[0] == '0') return false;
  }
  int max_window = 0, max_persist = 0;
  cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, L->ds->device);
  cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, L->ds->device);
  const size_t bytes = sizeof(short) * static_cast<size_t>(L->ds->num_data);
  if (max_window <= 0 || max_persist <= 0) return false;
  const size_t win = std::min(bytes, static_cast<size_t>(max_window));
  w->base_ptr = L->gh_q;
  w->num_bytes = win;
  // more bytes than may persist: pin the fraction that fits
  w->hitRatio = static_cast<float>(std::min(1.0, static_cast<double>(max_persist) / static_cast<double>(win)));
  w->hitProp = cudaAccessPropertyPersisting;
  w->missProp = cudaAccessPropertyStreaming;
  return true;
}

static void apply_gh_window_to_stream(Learner* L) {
  cudaAccessPolicyWindow w{};
  cudaStreamAttrValue v{};
  if (gh_window(L, &w)) {
    int max_persist = 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, L->ds->device);
    size_t cur = 0;
    cudaDeviceGetLimit(&cur, cudaLimitPersistingL2CacheSize);
    const size_t want = std::min(static_cast<size_t>(max_persist), w.num_bytes);
    if (cur < want) cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want);
    v.accessPolicyWindow = w;
  } else {
    v.accessPolicyWindow.num_bytes = 0;  // no window
  }
  cudaStreamSetAttribute(L->stream, cudaStreamAttributeAccessPolicyWindow, &v);
  cudaGetLastError();  // a device without the feature keeps working without it
}

static void apply_gh_window_to_graph(Learner* L, cudaGraph_t g) {
  cudaAccessPolicyWindow w{};
  if (!gh_window(L, &w)) return;
  size_t n = 0;
  if (cudaGraphGetNodes(g, nullptr, &n) != cudaSuccess || n == 0) return;
  std::vector<cudaGraphNode_t> nodes(n);
  if (cudaGraphGetNodes(g, nodes.data(), &n) != cudaSuccess) return;
  cudaKernelNodeAttrValue v{};
  v.accessPolicyWindow = w;
  for (cudaGraphNode_t nd : nodes) {
    cudaGraphNodeType ty;
    if (cudaGraphNodeGetType(nd, &ty) == cudaSuccess && ty == cudaGraphNodeTypeKernel)
      cudaGraphKernelNodeSetAttribute(nd, cudaKernelNodeAttributeAccessPolicyWindow, &v);
  }
  cudaGetLastError();
}

// ---- the three parts of a tree.  prologue: root histogram, root scan, first wave selected.  wave: split every
// node of the wave, histograms of the smaller children, scan of all children, next wave selected.  epilogue: replay.
static void enqueue_prologue(Learner* L, cudaStream_t s, cudaGraphConditionalHandle cond, int use_cond) {
  const LearnerDev& d = L->dev;
  const bool multi = L->comm != nullptr && L->comm->num_ranks > 1;
  {
    StageTimer tm(L, s, kStageOther);
    GBL_CUDA(cudaMemsetAsync(L->hist_pool, 0, sizeof(long long) * d.slot_stride, s));  // slot 0; later slots: part_scatter
    launch_tree_begin(d, L->bag_indices, L->bag_count, L->bag_on_device ? L->bag_counts_dev : nullptr, s);
    launch_maxabs(d, s);
  }
  if (multi) {
    // GradientDiscretizer: Network::GlobalSyncUpByMax (gradient_discretizer.cpp:101-104); also used for the
    // fixed-point scale in float mode so that every rank shares one scale.
    StageTimer tm(L, s, kStageComm);
    if (L->peer) {
      launch_xchg_max(d, s);
    } else {
      comm_allreduce_f32(L->comm, &L->ts->max_abs_g, 2, RedOp::kMax, s);
    }
  }
  {
    StageTimer tm(L, s, kStageOther);
    launch_scales_and_convert(d, s);
  }
  {
    StageTimer tm(L, s, kStageHist);
    launch_histogram(d, nullptr, -1, nullptr, s);
  }
  if (multi) {
    StageTimer tm(L, s, kStageComm);
    if (L->peer) {
      launch_xchg_wave(d, 1, s);
    } else {
      comm_allreduce_i64(L->comm, L->hist_pool, static_cast<size_t>(d.slot_stride), RedOp::kSum, s);
    }
  }
  {
    StageTimer tm(L, s, kStageFind);
    launch_root_init(d, s);
    launch_find_best(d, 1, cond, use_cond, s);
  }
}

// A wave in two halves.  Front: split the wave's nodes and build the smaller children's histograms -- everything
// it needs is on the device.  Back: (data-parallel: all-reduce of the wave's fresh slots, whose range only the
// host-driven loop knows: `first_slot` / `num_slots`), scan of all children, selection of the next wave.
static void enqueue_wave_front(Learner* L, cudaStream_t s) {
  const LearnerDev& d = L->dev;
  {
    StageTimer tm(L, s, kStagePart);
    launch_partition(d, s);
  }
  {
    StageTimer tm(L, s, kStageHist);
    launch_histogram(d, nullptr, -1, nullptr, s);
  }
}

static void enqueue_wave_back(Learner* L, cudaStream_t s, cudaGraphConditionalHandle cond, int use_cond, int first_slot,
                              int num_slots) {
  const LearnerDev& d = L->dev;
  const bool multi = L->comm != nullptr && L->comm->num_ranks > 1;
  if (multi && L->peer) {
    // the same sum over NVLink peer memory, sized on the device (exchange.cu): no host involvement, graph-capturable
    StageTimer tm(L, s, kStageComm);
    launch_xchg_wave(d, 0, s);
  } else if (multi && num_slots > 0) {
    // the wave's histograms + (smaller, larger) row counts in one all-reduce; replaces Network::ReduceScatter of
    // data_parallel_tree_learner.cpp:285-297 and the count sync of :457-458
    StageTimer tm(L, s, kStageComm);
    comm_allreduce_i64(L->comm, L->hist_pool + static_cast<long long>(first_slot) * d.slot_stride,
                       static_cast<size_t>(d.slot_stride) * num_slots, RedOp::kSum, s);
  }
  {
    StageTimer tm(L, s, kStageFind);
    launch_find_best(d, 0, cond, use_cond, s);
  }
}

static void enqueue_epilogue(Learner* L, cudaStream_t s) {
  const LearnerDev& d = L->dev;
  {
    StageTimer tm(L, s, kStageOther);
    launch_finalize(d, L->replay_scratch, s);
    if (L->cfg.use_quantized_grad && L->cfg.quant_train_renew_leaf) {
      // GradientDiscretizer::RenewIntGradTreeOutput (serial_tree_learner.cpp:238-241); data-parallel: GlobalSum of
      // the per-leaf sums (gradient_discretizer.cpp:231)
      launch_renew_sums(d, reinterpret_cast<double*>(L->unit_bits), L->renew_sums, s);
      if (L->comm != nullptr && L->comm->num_ranks > 1)
        comm_allreduce_f64(L->comm, L->renew_sums, 2 * static_cast<size_t>(L->cfg.num_leaves), RedOp::kSum, s);
      launch_renew_outputs(d, L->renew_sums, s);
    }
  }
  const int rounds = L->cfg.num_leaves - 1;
  GBL_CUDA(cudaMemcpyAsync(L->records_pinned, L->records, sizeof(gbl_split_record) * rounds, cudaMemcpyDeviceToHost, s));
  GBL_CUDA(cudaMemcpyAsync(L->ts_pinned, L->ts, sizeof(TreeState), cudaMemcpyDeviceToHost, s));
}

// Host-driven loop (profiling and the data-parallel learner): the host reads the tree state after every wave.
static void run_tree_host_loop(Learner* L, cudaStream_t s) {
  L->ev_used = 0;
  if (L->state_ready == nullptr) GBL_CUDA(cudaEventCreateWithFlags(&L->state_ready, cudaEventDisableTiming));
  enqueue_prologue(L, s, 0, 0);
  GBL_CUDA(cudaMemcpyAsync(L->ts_pinned, L->ts, sizeof(TreeState), cudaMemcpyDeviceToHost, s));
  GBL_CUDA(cudaEventRecord(L->state_ready, s));
  for (;;) {
    // the front half needs nothing from the host: enqueue it before waiting for the state of the previous
    // selection, so the GPU keeps working while the host reads it (a finished tree turns these launches into no-ops)
    enqueue_wave_front(L, s);
    GBL_CUDA(cudaEventSynchronize(L->state_ready));
    const TreeState t = *L->ts_pinned;
    if (std::getenv("GBL_DEBUG_WAVES") != nullptr)
      std::fprintf(stderr, "[wave %d] cand %d replayed %d expanded %d nodes %d last %d done %d\n", t.waves, t.num_cand, t.rp_ns,
                   t.num_expanded, t.num_nodes, t.last_wave, t.done);
    if (t.done) break;
    if (t.num_cand <= 0 || t.num_cand > L->dev.max_cand || t.num_slots + t.num_cand > L->dev.max_slots)
      fatal("learner: corrupt wave state (num_cand=%d, num_slots=%d)", t.num_cand, t.num_slots);
    enqueue_wave_back(L, s, 0, 0, t.num_slots, t.last_wave ? 0 : t.num_cand);
    GBL_CUDA(cudaMemcpyAsync(L->ts_pinned, L->ts, sizeof(TreeState), cudaMemcpyDeviceToHost, s));
    GBL_CUDA(cudaEventRecord(L->state_ready, s));
  }
  enqueue_epilogue(L, s);
}

// One CUDA graph per learner: prologue -> WHILE(conditional node; body = one wave) -> epilogue.  The scan kernel
// that selects the next wave also sets the loop condition, so a whole tree runs without the host.
static void build_graph(Learner* L) {
  cudaStream_t s = L->stream;
  cudaGraph_t graph = nullptr;
  GBL_CUDA(cudaGraphCreate(&graph, 0));
  try {
    cudaGraphConditionalHandle cond;
    GBL_CUDA(cudaGraphConditionalHandleCreate(&cond, graph, 0, cudaGraphCondAssignDefault));
    cudaGraph_t got = nullptr;
    // prologue
    GBL_CUDA(cudaStreamBeginCaptureToGraph(s, graph, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
    enqueue_prologue(L, s, cond, 1);
    cudaStreamCaptureStatus st;
    const cudaGraphNode_t* deps = nullptr;
    size_t ndeps = 0;
    GBL_CUDA(cudaStreamGetCaptureInfo_v2(s, &st, nullptr, nullptr, &deps, &ndeps));
    std::vector<cudaGraphNode_t> tail(deps, deps + ndeps);
    GBL_CUDA(cudaStreamEndCapture(s, &got));
    // while node
    cudaGraphNodeParams cp = {};
    cp.type = cudaGraphNodeTypeConditional;
    cp.conditional.handle = cond;
    cp.conditional.type = cudaGraphCondTypeWhile;
    cp.conditional.size = 1;
    cudaGraphNode_t loop = nullptr;
    GBL_CUDA(cudaGraphAddNode(&loop, graph, tail.data(), tail.size(), &cp));
    cudaGraph_t body = cp.conditional.phGraph_out[0];
    GBL_CUDA(cudaStreamBeginCaptureToGraph(s, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
    enqueue_wave_front(L, s);
    enqueue_wave_back(L, s, cond, 1, 0, 0);
    GBL_CUDA(cudaStreamEndCapture(s, &got));
    // epilogue
    GBL_CUDA(cudaStreamBeginCaptureToGraph(s, graph, &loop, nullptr, 1, cudaStreamCaptureModeThreadLocal));
    enqueue_epilogue(L, s);
    GBL_CUDA(cudaStreamEndCapture(s, &got));
    apply_gh_window_to_graph(L, graph);
    apply_gh_window_to_graph(L, body);
    GBL_CUDA(cudaGraphInstantiate(&L->graph_exec, graph, 0));
    L->graph_cond = cond;
  } catch (...) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(s, &st);
    if (st != cudaStreamCaptureStatusNone) {
      cudaGraph_t junk = nullptr;
      cudaStreamEndCapture(s, &junk);
    }
    cudaGraphDestroy(graph);
    throw;
  }
  cudaGraphDestroy(graph);
}

static void train_async(Learner* L, const float* grad_dev, const float* hess_dev) {
  if (grad_dev == nullptr || hess_dev == nullptr) fatal("gbl_learner_train: null gradients");
  if (L->cfg.use_quantized_grad && L->cfg.stochastic_rounding && L->rand_g == nullptr)
    fatal("learner: stochastic_rounding=1 needs gbl_learner_set_quant_random (host mt19937 arrays of the reference), "
          "or set stochastic_rounding=0");
  DeviceGuard guard(L->ds->device);
  LearnerDev& d = L->dev;
  d.grad = grad_dev;
  d.hess = hess_dev;
  d.feature_mask = L->feature_mask;
  d.rand_g = (L->cfg.use_quantized_grad && L->cfg.stochastic_rounding) ? L->rand_g : nullptr;
  d.rand_h = d.rand_g != nullptr ? L->rand_h : nullptr;
  d.rand_start = L->rand_start;
  const bool multi = L->comm != nullptr && L->comm->num_ranks > 1;
  if (L->use_graph && (!multi || L->peer) && !L->profiling) {
    const bool stale = L->graph_exec == nullptr || L->graph_grad != grad_dev || L->graph_hess != hess_dev ||
                       L->graph_rand_start != d.rand_start || L->graph_bag != L->bag_indices ||
                       L->graph_bag_count != (L->bag_on_device ? -2 : L->bag_count);
    if (stale) {
      drop_graph(L);
      const long long launches_before = g_kernel_launches.load();
      build_graph(L);
      g_kernel_launches.store(launches_before);  // launches "made" during capture only built graph nodes
      L->graph_grad = grad_dev;
      L->graph_hess = hess_dev;
      L->graph_rand_start = d.rand_start;
      L->graph_bag = L->bag_indices;
      L->graph_bag_count = L->bag_on_device ? -2 : L->bag_count;  // device-drawn bags: the size is read on the device
    }
    GBL_CUDA(cudaGraphLaunch(L->graph_exec, L->stream));
    L->tree_in_graph = true;
  } else {
    run_tree_host_loop(L, L->stream);
    L->tree_in_graph = false;
  }
  L->tree_pending = true;
  L->has_tree = true;
}

// kernels a tree launched, derived from the device's own counters (the graph replays a data-dependent number of
// waves): prologue (tree_begin, [bagged index copy], maxabs, scales, histogram, root_init, scan, select) + 6 per
// wave + the replay
static long long kernels_of_tree(const Learner* L, const TreeState& t) {
  const long long per_xchg = xchg_launches_per_wave();  // barrier / reduce / barrier, or the two- / one-kernel forms
  const long long xchg = L->peer ? 1 + per_xchg * (1 + t.waves) : 0;  // max exchange, root sum, one exchange per wave
  return (L->bag_indices != nullptr ? 8 : 7) + 6LL * t.waves + 1 + xchg +
         ((L->cfg.use_quantized_grad && L->cfg.quant_train_renew_leaf) ? 3 : 0);
}

static void train_finish(Learner* L, gbl_tree_result* result, gbl_split_record* records) {
  if (!L->tree_pending) fatal("gbl_learner_train_finish: no tree in flight");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  L->tree_pending = false;
  const TreeState& t = *L->ts_pinned;
  const int ns = t.num_splits;
  if (ns < 0 || ns > L->cfg.num_leaves - 1) fatal("learner: corrupt tree state (num_splits=%d)", ns);
  if (t.error != 0)
    fatal("learner: internal error %d while growing the tree (%s)", t.error,
          t.error == 1 ? "the replay needed a node that was never expanded"
                       : t.error == 2 ? "expansion capacity exhausted"
                                      : t.error == 9 ? "a rank did not reach a peer-memory barrier in time" : "a scan job was not reduced");
  if (L->tree_in_graph) g_kernel_launches.fetch_add(kernels_of_tree(L, t), std::memory_order_relaxed);
  if (L->profiling) {
    for (size_t i = 0; i + 1 < L->ev_used; i += 2) {
      float ms = 0.f;
      GBL_CUDA(cudaEventElapsedTime(&ms, L->ev[i], L->ev[i + 1]));
      switch (L->ev_stage[i / 2]) {
        case kStageHist: L->prof.hist_ms += ms; break;
        case kStageFind: L->prof.find_ms += ms; break;
        case kStagePart: L->prof.partition_ms += ms; break;
        case kStageComm: L->prof.collective_ms += ms; break;
        default: L->prof.other_ms += ms; break;
      }
    }
    L->prof.trees += 1;
  }
  L->prof.hist_launches += t.hist_launches;  // counted on the device whether or not stages are timed
  L->prof.hist_rows += t.hist_rows_total;
  L->prof.partition_rows += t.part_rows_total;
  L->prof.waves += t.waves;
  L->prof.expansions += t.num_expanded;
  L->prof.splits += ns;
  const bool multi = L->comm != nullptr && L->comm->num_ranks > 1;
  if (multi && ns > 0) {
    // global leaf counts of every split (data_parallel_tree_learner.cpp:457-458): one all-reduce per tree
    std::vector<long long> counts(2 * ns);
    for (int k = 0; k < ns; ++k) {
      counts[2 * k] = L->records_pinned[k].left_count;
      counts[2 * k + 1] = L->records_pinned[k].right_count;
    }
    long long* buf = reinterpret_cast<long long*>(L->cand);  // >= 2*F*sizeof(BestSplit) bytes; reuse when it fits
    long long* tmp = nullptr;
    if (sizeof(long long) * 2 * ns > sizeof(BestSplit) * 2 * L->ds->num_features) {
      GBL_CUDA(cudaMalloc(&tmp, sizeof(long long) * 2 * ns));
      buf = tmp;
    }
    GBL_CUDA(cudaMemcpyAsync(buf, counts.data(), sizeof(long long) * 2 * ns, cudaMemcpyHostToDevice, L->stream));
    comm_allreduce_i64(L->comm, buf, 2 * ns, RedOp::kSum, L->stream);
    GBL_CUDA(cudaMemcpyAsync(counts.data(), buf, sizeof(long long) * 2 * ns, cudaMemcpyDeviceToHost, L->stream));
    GBL_CUDA(cudaStreamSynchronize(L->stream));
    if (tmp) cudaFree(tmp);
    for (int k = 0; k < ns; ++k) {
      L->records_pinned[k].left_count = static_cast<int>(counts[2 * k]);
      L->records_pinned[k].right_count = static_cast<int>(counts[2 * k + 1]);
    }
  }
  if (result != nullptr) {
    std::memset(result, 0, sizeof(*result));
    result->num_splits = ns;
    result->root_output = t.root_output;
    result->grad_scale = t.q_grad_scale;
    result->hess_scale = t.q_hess_scale;
  }
  if (records != nullptr && ns > 0) std::memcpy(records, L->records_pinned, sizeof(gbl_split_record) * ns);
}

}  // namespace gbl

using namespace gbl;

#define LEARNER(h) (reinterpret_cast<Learner*>(h))

extern "C" {

int gbl_learner_create(const gbl_dataset* ds, const gbl_learner_config* config, gbl_comm* comm, gbl_learner** out) {
  GBL_API_BEGIN();
  if (out == nullptr) fatal("gbl_learner_create: null argument");
  *out = reinterpret_cast<gbl_learner*>(
      learner_create(reinterpret_cast<const Dataset*>(ds), config, reinterpret_cast<Comm*>(comm)));
  GBL_API_END();
}

int gbl_learner_free(gbl_learner* learner) {
  GBL_API_BEGIN();
  learner_free(LEARNER(learner));
  GBL_API_END();
}

int gbl_learner_reset_config(gbl_learner* learner, const gbl_learner_config* config) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || config == nullptr) fatal("gbl_learner_reset_config: null argument");
  validate_config(*config);
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  drop_graph(L);
  const bool realloc = config->num_leaves != L->cfg.num_leaves || config->use_quantized_grad != L->cfg.use_quantized_grad;
  L->cfg = *config;
  teardown_peer(L);
  if (realloc) {
    free_tree_buffers(L);
    alloc_tree_buffers(L);
    build_records(L);
    L->has_tree = false;
  } else {
    L->dev.p = make_params(L);
  }
  setup_peer(L);  // collective: the ranks of a data-parallel learner reset their configuration together
  GBL_API_END();
}

int gbl_learner_set_monotone(gbl_learner* learner, int32_t enabled, double monotone_penalty) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_set_monotone: null learner");
  if (monotone_penalty < 0.0) fatal("gbl_learner_set_monotone: monotone_penalty must be >= 0");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  drop_graph(L);  // the kernels take the parameters by value
  L->use_mc = enabled != 0;
  L->monotone_penalty = monotone_penalty;
  L->dev.p = make_params(L);
  GBL_API_END();
}

int gbl_learner_set_categorical(gbl_learner* learner, int32_t max_cat_threshold, double cat_l2, double cat_smooth,
                                int32_t max_cat_to_onehot, int32_t min_data_per_group) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_set_categorical: null learner");
  if (max_cat_threshold <= 0) fatal("gbl_learner_set_categorical: max_cat_threshold must be > 0");
  if (cat_l2 < 0.0 || cat_smooth < 0.0) fatal("gbl_learner_set_categorical: cat_l2 / cat_smooth must be >= 0");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  drop_graph(L);
  L->max_cat_threshold = max_cat_threshold;
  L->cat_l2 = cat_l2;
  L->cat_smooth = cat_smooth;
  L->max_cat_to_onehot = max_cat_to_onehot;
  L->min_data_per_group = min_data_per_group;
  L->dev.p = make_params(L);
  GBL_API_END();
}

int gbl_learner_set_feature_mask(gbl_learner* learner, const int8_t* is_feature_used_host) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_set_feature_mask: null learner");
  DeviceGuard guard(L->ds->device);
  if (is_feature_used_host != nullptr && L->feature_mask != nullptr) {
    // a new mask in the same buffer (ColSampler::ResetByTree before every tree): ordered behind the previous tree on the
    // learner's stream, and the captured tree stays valid because it reads the mask through the same pointer
    GBL_CUDA(cudaMemcpyAsync(L->feature_mask, is_feature_used_host, L->ds->num_features, cudaMemcpyHostToDevice, L->stream));
  } else {
    GBL_CUDA(cudaStreamSynchronize(L->stream));
    drop_graph(L);
    if (is_feature_used_host == nullptr) {
      cudaFree(L->feature_mask);
      L->feature_mask = nullptr;
    } else {
      GBL_CUDA(cudaMalloc(&L->feature_mask, L->ds->num_features));
      GBL_CUDA(cudaMemcpy(L->feature_mask, is_feature_used_host, L->ds->num_features, cudaMemcpyHostToDevice));
    }
  }
  GBL_API_END();
}

int gbl_learner_set_bagging_data(gbl_learner* learner, const int32_t* used_indices_host, int32_t num_used) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_set_bagging_data: null learner");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  drop_graph(L);
  L->bag_on_device = false;
  if (used_indices_host == nullptr || num_used <= 0 || num_used >= L->ds->num_data) {
    cudaFree(L->bag_indices);
    cudaFree(L->oob_indices);
    L->bag_indices = nullptr;
    L->oob_indices = nullptr;
    L->bag_count = 0;
    L->oob_count = 0;
  } else {
    if (L->comm != nullptr && L->comm->num_ranks > 1) fatal("learner: bagging with the data-parallel learner is not supported");
    if (L->cfg.use_quantized_grad) fatal("learner: bagging with use_quantized_grad is not supported");
    if (L->bag_indices == nullptr) GBL_CUDA(cudaMalloc(&L->bag_indices, sizeof(int) * L->ds->num_data));
    GBL_CUDA(cudaMemcpy(L->bag_indices, used_indices_host, sizeof(int) * num_used, cudaMemcpyHostToDevice));
    L->bag_count = num_used;
    // the complement, for the score update of the rows the tree is not trained on
    const int n = static_cast<int>(L->ds->num_data);
    std::vector<char> in_bag(n, 0);
    for (int i = 0; i < num_used; ++i) {
      const int r = used_indices_host[i];
      if (r < 0 || r >= n || in_bag[r]) fatal("gbl_learner_set_bagging_data: indices must be distinct rows of the dataset");
      in_bag[r] = 1;
    }
    std::vector<int> oob;
    oob.reserve(n - num_used);
    for (int r = 0; r < n; ++r) {
      if (!in_bag[r]) oob.push_back(r);
    }
    if (L->oob_indices == nullptr) GBL_CUDA(cudaMalloc(&L->oob_indices, sizeof(int) * n));
    GBL_CUDA(cudaMemcpy(L->oob_indices, oob.data(), sizeof(int) * oob.size(), cudaMemcpyHostToDevice));
    L->oob_count = static_cast<int>(oob.size());
  }
  GBL_API_END();
}

static void bagging_dev(Learner* L, int32_t bagging_seed, double fraction, bool balanced, double pos_fraction,
                        double neg_fraction, const float* label_host, int32_t reset, int32_t* out_bag_count,
                        const int32_t* query_boundaries_host = nullptr, int32_t num_queries = 0, bool by_query = false) {
  auto in_unit = [](double f) { return f > 0.0 && f <= 1.0; };
  if (!balanced && !in_unit(fraction)) fatal("gbl_learner_bagging_dev: bagging_fraction must be in (0, 1]");
  if (balanced && (!in_unit(pos_fraction) || !in_unit(neg_fraction)))
    fatal("gbl_learner_bagging_balanced_dev: pos_bagging_fraction / neg_bagging_fraction must be in (0, 1]");
  if (L->comm != nullptr && L->comm->num_ranks > 1) fatal("learner: bagging with the data-parallel learner is not supported");
  if (L->cfg.use_quantized_grad) fatal("learner: bagging with use_quantized_grad is not supported");
  DeviceGuard guard(L->ds->device);
  const int n = static_cast<int>(L->ds->num_data);
  const int nblk = (n + 1023) / 1024;
  if (balanced) {
    if (label_host != nullptr) {
      if (L->bag_label == nullptr) GBL_CUDA(cudaMalloc(&L->bag_label, sizeof(float) * n));
      GBL_CUDA(cudaMemcpyAsync(L->bag_label, label_host, sizeof(float) * n, cudaMemcpyHostToDevice, L->stream));
    } else if (L->bag_label == nullptr) {
      fatal("gbl_learner_bagging_balanced_dev: the first call needs the labels");
    }
  }
  if (by_query) {
    if (query_boundaries_host != nullptr) {
      if (num_queries <= 0 || query_boundaries_host[0] != 0 || query_boundaries_host[num_queries] != n)
        fatal("gbl_learner_bagging_by_query_dev: query boundaries must cover rows 0..num_data");
      std::vector<int> row_query(n);
      for (int q = 0; q < num_queries; ++q) {
        if (query_boundaries_host[q + 1] < query_boundaries_host[q]) fatal("gbl_learner_bagging_by_query_dev: boundaries must not decrease");
        for (int r = query_boundaries_host[q]; r < query_boundaries_host[q + 1]; ++r) row_query[r] = q;
      }
      cudaFree(L->bag_row_query);
      cudaFree(L->bag_qbits);
      cudaFree(L->bag_qblk_cnt);
      L->bag_row_query = nullptr;
      const int nblk_q = (num_queries + 1023) / 1024;
      GBL_CUDA(cudaMalloc(&L->bag_row_query, sizeof(int) * n));
      GBL_CUDA(cudaMalloc(&L->bag_qbits, sizeof(unsigned) * nblk_q * 32));
      GBL_CUDA(cudaMalloc(&L->bag_qblk_cnt, sizeof(int) * nblk_q));
      GBL_CUDA(cudaMemcpy(L->bag_row_query, row_query.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
      L->bag_num_queries = num_queries;
    } else if (L->bag_row_query == nullptr) {
      fatal("gbl_learner_bagging_by_query_dev: the first call needs the query boundaries");
    }
  }
  if (L->bag_state == nullptr) {
    GBL_CUDA(cudaMalloc(&L->bag_state, sizeof(unsigned) * nblk));
    GBL_CUDA(cudaMalloc(&L->bag_bits, sizeof(unsigned) * nblk * 32));
    GBL_CUDA(cudaMalloc(&L->bag_blk_cnt, sizeof(int) * nblk));
    GBL_CUDA(cudaMalloc(&L->bag_blk_off, sizeof(int) * nblk));
    GBL_CUDA(cudaMalloc(&L->bag_counts_dev, sizeof(int) * 2));
    GBL_CUDA(cudaMallocHost(&L->bag_counts_pinned, sizeof(int) * 2));
    reset = 1;
  }
  if (!L->bag_on_device || L->bag_indices == nullptr || L->oob_indices == nullptr) {
    // first device-drawn bag (or after a host-provided one): the captured tree must read the bag size on the device
    GBL_CUDA(cudaStreamSynchronize(L->stream));
    drop_graph(L);
    if (L->bag_indices == nullptr) GBL_CUDA(cudaMalloc(&L->bag_indices, sizeof(int) * n));
    if (L->oob_indices == nullptr) GBL_CUDA(cudaMalloc(&L->oob_indices, sizeof(int) * n));
    L->bag_on_device = true;
  }
  if (reset || bagging_seed != L->bag_seed) {
    launch_bag_seed(L->bag_state, nblk, bagging_seed, L->stream);
    L->bag_seed = bagging_seed;
  }
  if (by_query) {
    // the generators are indexed by QUERY block (bagging.hpp:36-66 runs BaggingHelper over num_queries_ items)
    launch_bag_draw_by_query(L->bag_state, (L->bag_num_queries + 1023) / 1024, L->bag_num_queries, fraction, L->bag_qbits,
                             L->bag_qblk_cnt, L->bag_row_query, nblk, n, L->bag_bits, L->bag_blk_cnt, L->bag_blk_off,
                             L->bag_counts_dev, L->bag_indices, L->oob_indices, L->stream);
  } else {
    launch_bag_draw(L->bag_state, nblk, n, fraction, balanced ? L->bag_label : nullptr, pos_fraction, neg_fraction, L->bag_bits,
                    L->bag_blk_cnt, L->bag_blk_off, L->bag_counts_dev, L->bag_indices, L->oob_indices, L->stream);
  }
  GBL_CUDA(cudaMemcpyAsync(L->bag_counts_pinned, L->bag_counts_dev, sizeof(int) * 2, cudaMemcpyDeviceToHost, L->stream));
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  L->bag_count = L->bag_counts_pinned[0];
  L->oob_count = L->bag_counts_pinned[1];
  if (out_bag_count != nullptr) *out_bag_count = L->bag_count;
}

int gbl_learner_bagging_dev(gbl_learner* learner, int32_t bagging_seed, double bagging_fraction, int32_t reset,
                            int32_t* out_bag_count) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_bagging_dev: null learner");
  bagging_dev(L, bagging_seed, bagging_fraction, false, 1.0, 1.0, nullptr, reset, out_bag_count);
  GBL_API_END();
}

int gbl_learner_bagging_balanced_dev(gbl_learner* learner, int32_t bagging_seed, double pos_bagging_fraction,
                                     double neg_bagging_fraction, const float* label_host, int32_t reset,
                                     int32_t* out_bag_count) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_bagging_balanced_dev: null learner");
  bagging_dev(L, bagging_seed, 1.0, true, pos_bagging_fraction, neg_bagging_fraction, label_host, reset, out_bag_count);
  GBL_API_END();
}

int gbl_learner_bagging_by_query_dev(gbl_learner* learner, int32_t bagging_seed, double bagging_fraction,
                                     const int32_t* query_boundaries_host, int32_t num_queries, int32_t reset,
                                     int32_t* out_bag_count) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_bagging_by_query_dev: null learner");
  bagging_dev(L, bagging_seed, bagging_fraction, false, 1.0, 1.0, nullptr, reset, out_bag_count, query_boundaries_host,
              num_queries, true);
  GBL_API_END();
}

int gbl_learner_set_quant_random(gbl_learner* learner, const double* grad_random_host, const double* hess_random_host,
                                 int64_t n) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || grad_random_host == nullptr || hess_random_host == nullptr) fatal("gbl_learner_set_quant_random: null argument");
  if (n != L->ds->num_data) fatal("gbl_learner_set_quant_random: need num_data=%lld values", (long long)L->ds->num_data);
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  drop_graph(L);
  if (L->rand_g == nullptr) {
    GBL_CUDA(cudaMalloc(&L->rand_g, sizeof(double) * n));
    GBL_CUDA(cudaMalloc(&L->rand_h, sizeof(double) * n));
  }
  GBL_CUDA(cudaMemcpy(L->rand_g, grad_random_host, sizeof(double) * n, cudaMemcpyHostToDevice));
  GBL_CUDA(cudaMemcpy(L->rand_h, hess_random_host, sizeof(double) * n, cudaMemcpyHostToDevice));
  L->rand_n = n;
  GBL_API_END();
}

int gbl_learner_set_quant_random_start(gbl_learner* learner, int32_t random_values_use_start) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_set_quant_random_start: null learner");
  L->rand_start = random_values_use_start;
  GBL_API_END();
}

int gbl_learner_train_async(gbl_learner* learner, const float* gradients_dev, const float* hessians_dev) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_train_async: null learner");
  if (L->tree_pending) {
    DeviceGuard guard(L->ds->device);
    GBL_CUDA(cudaStreamSynchronize(L->stream));
    L->tree_pending = false;
  }
  train_async(L, gradients_dev, hessians_dev);
  GBL_API_END();
}

int gbl_learner_train_finish(gbl_learner* learner, gbl_tree_result* result_host, gbl_split_record* records_host) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_train_finish: null learner");
  train_finish(L, result_host, records_host);
  GBL_API_END();
}

int gbl_learner_train(gbl_learner* learner, const float* gradients_dev, const float* hessians_dev,
                      gbl_tree_result* result_host, gbl_split_record* records_host) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_train: null learner");
  train_async(L, gradients_dev, hessians_dev);
  train_finish(L, result_host, records_host);
  GBL_API_END();
}

int gbl_learner_train_host(gbl_learner* learner, const float* gradients_host, const float* hessians_host,
                           gbl_tree_result* result_host, gbl_split_record* records_host) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || gradients_host == nullptr || hessians_host == nullptr) fatal("gbl_learner_train_host: null argument");
  DeviceGuard guard(L->ds->device);
  const size_t n = static_cast<size_t>(L->ds->num_data);
  if (L->grad_stage == nullptr) {
    GBL_CUDA(cudaMalloc(&L->grad_stage, sizeof(float) * n));
    GBL_CUDA(cudaMalloc(&L->hess_stage, sizeof(float) * n));
  }
  GBL_CUDA(cudaMemcpyAsync(L->grad_stage, gradients_host, sizeof(float) * n, cudaMemcpyHostToDevice, L->stream));
  GBL_CUDA(cudaMemcpyAsync(L->hess_stage, hessians_host, sizeof(float) * n, cudaMemcpyHostToDevice, L->stream));
  train_async(L, L->grad_stage, L->hess_stage);
  train_finish(L, result_host, records_host);
  GBL_API_END();
}

int gbl_learner_add_prediction_to_score(gbl_learner* learner, const double* leaf_values_host, int32_t num_leaves,
                                        double* score_dev) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || leaf_values_host == nullptr || score_dev == nullptr) fatal("gbl_learner_add_prediction_to_score: null argument");
  if (!L->has_tree) fatal("gbl_learner_add_prediction_to_score: no tree trained yet");
  if (num_leaves < 1 || num_leaves > L->cfg.num_leaves) fatal("gbl_learner_add_prediction_to_score: bad num_leaves");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaMemcpyAsync(L->leaf_values_tmp, leaf_values_host, sizeof(double) * num_leaves, cudaMemcpyHostToDevice, L->stream));
  launch_add_score(L->dev, L->leaf_values_tmp, 1.0, 0.0, score_dev, L->leaf_u16, L->bag_indices != nullptr, L->stream);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  GBL_API_END();
}

int gbl_learner_add_prediction_to_score_dev(gbl_learner* learner, double shrinkage, double bias, double* score_dev) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || score_dev == nullptr) fatal("gbl_learner_add_prediction_to_score_dev: null argument");
  if (!L->has_tree) fatal("gbl_learner_add_prediction_to_score_dev: no tree trained yet");
  DeviceGuard guard(L->ds->device);
  launch_add_score(L->dev, L->leaf_values, shrinkage, bias, score_dev, L->leaf_u16, L->bag_indices != nullptr, L->stream);
  GBL_API_END();
}

int gbl_learner_add_prediction_out_of_bag_dev(gbl_learner* learner, double shrinkage, double bias, double* score_dev) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || score_dev == nullptr) fatal("gbl_learner_add_prediction_out_of_bag_dev: null argument");
  if (!L->has_tree) fatal("gbl_learner_add_prediction_out_of_bag_dev: no tree trained yet");
  DeviceGuard guard(L->ds->device);
  if (L->bag_indices != nullptr && L->oob_count > 0) {
    launch_add_score_rows(L->dev, L->ds->row_bins, L->ds->row_stride, L->oob_indices, L->oob_count, L->leaf_values, shrinkage,
                          bias, score_dev, L->stream);
  }
  GBL_API_END();
}

int gbl_learner_add_prediction_to_dataset_dev(gbl_learner* learner, const gbl_dataset* other, double shrinkage, double bias,
                                              double* score_dev) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  const Dataset* o = reinterpret_cast<const Dataset*>(other);
  if (L == nullptr || o == nullptr || score_dev == nullptr) fatal("gbl_learner_add_prediction_to_dataset_dev: null argument");
  if (!L->has_tree) fatal("gbl_learner_add_prediction_to_dataset_dev: no tree trained yet");
  if (o->device != L->ds->device) fatal("gbl_learner_add_prediction_to_dataset_dev: datasets on different devices");
  if (o->num_features != L->ds->num_features)
    fatal("gbl_learner_add_prediction_to_dataset_dev: %d features, the training data has %d (same bin mappers required)",
          o->num_features, L->ds->num_features);
  DeviceGuard guard(L->ds->device);
  launch_add_score_rows(L->dev, o->row_bins, o->row_stride, nullptr, static_cast<int>(o->num_data), L->leaf_values, shrinkage,
                        bias, score_dev, L->stream);
  GBL_API_END();
}

int gbl_learner_get_leaf_values(gbl_learner* learner, double* out_host, int32_t num_leaves) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_host == nullptr) fatal("gbl_learner_get_leaf_values: null argument");
  if (!L->has_tree) fatal("gbl_learner_get_leaf_values: no tree trained yet");
  if (num_leaves < 1 || num_leaves > L->cfg.num_leaves) fatal("gbl_learner_get_leaf_values: bad num_leaves");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  GBL_CUDA(cudaMemcpy(out_host, L->leaf_values, sizeof(double) * num_leaves, cudaMemcpyDeviceToHost));
  GBL_API_END();
}

int gbl_learner_get_leaf_assignment(gbl_learner* learner, int32_t* out_host) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_host == nullptr) fatal("gbl_learner_get_leaf_assignment: null argument");
  if (!L->has_tree) fatal("gbl_learner_get_leaf_assignment: no tree trained yet");
  DeviceGuard guard(L->ds->device);
  const size_t n = static_cast<size_t>(L->ds->num_data);
  int* tmp = nullptr;
  GBL_CUDA(cudaMalloc(&tmp, sizeof(int) * n));
  GBL_CUDA(cudaMemsetAsync(tmp, 0xff, sizeof(int) * n, L->stream));
  launch_leaf_assignment(L->dev, tmp, L->stream);
  GBL_CUDA(cudaMemcpyAsync(out_host, tmp, sizeof(int) * n, cudaMemcpyDeviceToHost, L->stream));
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  cudaFree(tmp);
  GBL_API_END();
}

// rows of one final leaf: the union of the row segments below it, sorted (the reference's DataPartition keeps the
// rows of a leaf in ascending order; segments of a speculatively expanded leaf are each ascending)
static void download_nodes(Learner* L, std::vector<Node>* nodes, TreeState* ts) {
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  GBL_CUDA(cudaMemcpy(ts, L->ts, sizeof(TreeState), cudaMemcpyDeviceToHost));
  nodes->resize(ts->num_nodes);
  GBL_CUDA(cudaMemcpy(nodes->data(), L->nodes, sizeof(Node) * ts->num_nodes, cudaMemcpyDeviceToHost));
}

int gbl_learner_get_leaf_assignment_u16(gbl_learner* learner, uint16_t* out_host) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_host == nullptr) fatal("gbl_learner_get_leaf_assignment_u16: null argument");
  if (!L->has_tree) fatal("gbl_learner_get_leaf_assignment_u16: no tree trained yet");
  if (L->cfg.num_leaves > 65535) fatal("gbl_learner_get_leaf_assignment_u16: num_leaves=%d does not fit 16 bits", L->cfg.num_leaves);
  DeviceGuard guard(L->ds->device);
  const size_t n = static_cast<size_t>(L->ds->num_data);
  launch_leaf_mark(L->dev, L->leaf_u16, L->bag_indices != nullptr, L->stream);
  GBL_CUDA(cudaMemcpyAsync(out_host, L->leaf_u16, sizeof(unsigned short) * n, cudaMemcpyDeviceToHost, L->stream));
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  GBL_API_END();
}

int gbl_learner_get_leaf_indices(gbl_learner* learner, int32_t leaf, int32_t* out_host, int32_t* out_count) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_count == nullptr) fatal("gbl_learner_get_leaf_indices: null argument");
  if (!L->has_tree) fatal("gbl_learner_get_leaf_indices: no tree trained yet");
  if (leaf < 0 || leaf >= L->cfg.num_leaves) fatal("gbl_learner_get_leaf_indices: leaf out of range");
  DeviceGuard guard(L->ds->device);
  std::vector<Node> nodes;
  TreeState ts;
  download_nodes(L, &nodes, &ts);
  int total = 0;
  for (const Node& nd : nodes) {
    if (nd.left < 0 && nd.ref_leaf == leaf) total += nd.count;
  }
  *out_count = leaf <= ts.num_splits ? total : 0;
  if (out_host != nullptr && leaf <= ts.num_splits && total > 0) {
    int at = 0, segments = 0;
    for (const Node& nd : nodes) {
      if (nd.left < 0 && nd.ref_leaf == leaf && nd.count > 0) {
        GBL_CUDA(cudaMemcpy(out_host + at, L->idx[nd.buf] + nd.begin, sizeof(int) * nd.count, cudaMemcpyDeviceToHost));
        at += nd.count;
        ++segments;
      }
    }
    if (segments > 1) std::sort(out_host, out_host + total);
  }
  GBL_API_END();
}

// histogram of one final leaf.  A leaf that was expanded speculatively no longer owns a slot of its own (the larger
// child took it over): its histogram is the exact integer sum of the segments' histograms below it.
int gbl_learner_get_histogram(gbl_learner* learner, int32_t leaf, int32_t feature, double* out_host) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_host == nullptr) fatal("gbl_learner_get_histogram: null argument");
  if (!L->has_tree) fatal("gbl_learner_get_histogram: no tree trained yet");
  if (leaf < 0 || leaf >= L->cfg.num_leaves || feature < 0 || feature >= L->ds->num_features)
    fatal("gbl_learner_get_histogram: index out of range");
  DeviceGuard guard(L->ds->device);
  std::vector<Node> nodes;
  TreeState ts;
  download_nodes(L, &nodes, &ts);
  const long long words = L->dev.slot_stride;
  const int tile = feature / kTile, fl = feature % kTile;
  const int nb = L->ds->meta_host[feature].num_bin;
  const bool quant = L->cfg.use_quantized_grad != 0;
  std::vector<long long> slot(words);
  std::vector<long long> sum_g(nb, 0), sum_h(nb, 0);
  for (const Node& nd : nodes) {
    if (!(nd.left < 0 && nd.ref_leaf == leaf)) continue;
    GBL_CUDA(cudaMemcpy(slot.data(), L->hist_pool + static_cast<long long>(nd.slot) * words, sizeof(long long) * words,
                        cudaMemcpyDeviceToHost));
    for (int b = 0; b < nb; ++b) {
      const long long e = (static_cast<long long>(tile) * kMaxBin + b) * kTile + fl;
      if (quant) {
        const long long v = slot[e];
        sum_g[b] += static_cast<long long>(static_cast<int>(v >> 32));
        sum_h[b] += static_cast<long long>(static_cast<unsigned>(v & 0xffffffffLL));
      } else {
        sum_g[b] += slot[2 * e];
        sum_h[b] += slot[2 * e + 1];
      }
    }
  }
  for (int b = 0; b < nb; ++b) {
    out_host[2 * b] = quant ? static_cast<double>(sum_g[b]) : static_cast<double>(sum_g[b]) * ts.g_inv;
    out_host[2 * b + 1] = quant ? static_cast<double>(sum_h[b]) : static_cast<double>(sum_h[b]) * ts.h_inv;
  }
  GBL_API_END();
}

int gbl_learner_get_discretized(gbl_learner* learner, int8_t* out_host) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_host == nullptr) fatal("gbl_learner_get_discretized: null argument");
  if (!L->has_tree || !L->cfg.use_quantized_grad) fatal("gbl_learner_get_discretized: no quantized tree trained yet");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  GBL_CUDA(cudaMemcpy(out_host, L->gh_q, sizeof(short) * L->ds->num_data, cudaMemcpyDeviceToHost));
  GBL_API_END();
}

int gbl_learner_histogram_words(gbl_learner* learner, int64_t* out_words) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_words == nullptr) fatal("gbl_learner_histogram_words: null argument");
  *out_words = L->dev.slot_stride;
  GBL_API_END();
}

int gbl_learner_construct_histogram(gbl_learner* learner, const int32_t* indices_dev, int64_t num_indices,
                                    int64_t* out_hist_dev) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_hist_dev == nullptr) fatal("gbl_learner_construct_histogram: null argument");
  if (!L->has_tree) fatal("gbl_learner_construct_histogram: train one tree first (gradients are taken from it)");
  if (num_indices < 0 || num_indices > L->ds->num_data) fatal("gbl_learner_construct_histogram: bad num_indices");
  DeviceGuard guard(L->ds->device);
  launch_histogram(L->dev, indices_dev, static_cast<int>(num_indices), reinterpret_cast<long long*>(out_hist_dev),
                   L->stream);
  GBL_API_END();
}

int gbl_learner_set_profiling(gbl_learner* learner, int32_t enabled) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr) fatal("gbl_learner_set_profiling: null learner");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaStreamSynchronize(L->stream));
  drop_graph(L);
  L->profiling = enabled != 0;
  GBL_API_END();
}

int gbl_learner_timer_record(gbl_learner* learner, int32_t slot) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || slot < 0 || slot >= 8) fatal("gbl_learner_timer_record: bad argument");
  DeviceGuard guard(L->ds->device);
  if (L->timer[slot] == nullptr) GBL_CUDA(cudaEventCreate(&L->timer[slot]));
  GBL_CUDA(cudaEventRecord(L->timer[slot], L->stream));
  GBL_API_END();
}

int gbl_learner_timer_elapsed_ms(gbl_learner* learner, int32_t slot_begin, int32_t slot_end, double* out_ms) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out_ms == nullptr || slot_begin < 0 || slot_begin >= 8 || slot_end < 0 || slot_end >= 8 ||
      L->timer[slot_begin] == nullptr || L->timer[slot_end] == nullptr)
    fatal("gbl_learner_timer_elapsed_ms: bad argument");
  DeviceGuard guard(L->ds->device);
  GBL_CUDA(cudaEventSynchronize(L->timer[slot_end]));
  float ms = 0.f;
  GBL_CUDA(cudaEventElapsedTime(&ms, L->timer[slot_begin], L->timer[slot_end]));
  *out_ms = ms;
  GBL_API_END();
}

int gbl_learner_get_profile(gbl_learner* learner, gbl_profile* out, int32_t reset) {
  GBL_API_BEGIN();
  Learner* L = LEARNER(learner);
  if (L == nullptr || out == nullptr) fatal("gbl_learner_get_profile: null argument");
  *out = L->prof;
  if (reset) L->prof = gbl_profile{};
  GBL_API_END();
}

void* gbl_learner_stream(gbl_learner* learner) { return learner ? LEARNER(learner)->stream : nullptr; }

}  // extern "C"
