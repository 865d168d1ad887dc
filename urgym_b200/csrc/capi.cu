// C ABI of liburgym_b200.so (declared in include/urgym_b200.h). Host code only orchestrates: device memory, streams,
// launches. No CPU physics exists in this library: without a usable CUDA device every call fails.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <mutex>
#include <string>
#include <vector>

#if defined(UG_F64) && !defined(URGYM_F64)
#define URGYM_F64 1
#endif
#include "../../include/urgym_b200.h"
#ifndef UG_DEFAULT_GROUP
#define UG_DEFAULT_GROUP 32
#endif
#include "kernels.cuh"

// ---- urgym_step_static: the allocation-free entry (include/urgym_b200.h). Everything the launch needs besides the
// caller's buffers lives in ONE file-scope device array, which every GPU that loads this library holds until exit: the
// seven scene models (written once per device at set-up), the episode counters of the call (cleared by every call) and the
// finished-env flags that carry the auto-reset of the ball scenes from the step grid to the reset-only grid (every flag
// below num_envs is rewritten by the step grid before the reset grid reads it, so nothing there has to start at zero).
// In host code the array's name is only a stand-in for the symbol: its device address comes from cudaGetSymbolAddress
// with the device current, once per device.
#define UG_STATIC_MAX_ENVS 16384
struct StaticScratch {
  DevModel model[UG_ENV_COUNT];
  double stats[8];
  uint8_t defer[UG_STATIC_MAX_ENVS];
};
static __device__ __align__(256) unsigned char g_static_scratch[(sizeof(StaticScratch) + 255) / 256 * 256];

namespace {

thread_local std::string g_err;
std::atomic<long long> g_launches{0};

int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CK(call)                                                                                        \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) return fail(URGYM_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
  } while (0)

struct ModelCache { DevModel host; DevModel* dev = nullptr; int device = -1; };

}  // namespace

struct urgym_env_s {
  int env_type, N, device, gidx0, sparse;
  uint64_t seed;
  DevModel hmodel; DevModel* dmodel = nullptr;
  real* state = nullptr; int32_t* contacts = nullptr; int32_t* ccount = nullptr; uint32_t* hitmask = nullptr; double* stats = nullptr;
  // staging for the *_host entry points
  real *d_action = nullptr, *d_u = nullptr, *d_draws = nullptr, *d_obs = nullptr, *d_final = nullptr, *d_reward = nullptr;
  uint8_t *d_term = nullptr, *d_trunc = nullptr, *d_succ = nullptr, *d_mask = nullptr; int32_t* d_nticks = nullptr;
  uint8_t* d_defer = nullptr; bool split_reset = false;  // auto-reset in its own launch
  cudaStream_t own_stream = nullptr, last_stream = nullptr;
  cudaEvent_t order_evt = nullptr; bool used = false;   // cross-stream ordering of successive calls on one handle
};

namespace {

// One warp runs one env. Two envs per warp (16-lane groups) was measured in round 1 and lost on every
// scene; with up to 32 manifold points per tick the normal rows need the 32 lanes anyway.
int group_size() { return 32; }

// Whether auto-resets run in a launch of their own. Measured per env type (profiles/r01_reset_scheduling.txt): the
// ball scenes gain 7-19% because step ticks and the robot-only ticks of a reset no longer share an SM, the scenes whose
// reset settles for hundreds of ticks (Golf) lose to the latency of a thin second grid. URGYM_SPLIT_RESET=0/1 overrides.
// Round 2 (profiles/r02_split_vs_fused.txt): with the polytope gripper the ball stays on the paddle four times longer, so only
// ~5 % of the envs restart per step; at <= 4096 envs per GPU the step is then bound by the serial chain (step ticks + 162
// reset ticks of the slowest env) and the fused launch, which starts every reset the moment its env finishes, is 0-3 % ahead;
// from 8192 envs on the launch is throughput-bound and the split form wins by 6-20 %.
bool default_split_reset(int env_type, int num_envs) {
  static const bool dflt[UG_ENV_COUNT] = {true, true, false, false, false, false, false, false};
  const char* e = getenv("URGYM_SPLIT_RESET");
  return e ? atoi(e) != 0 : (dflt[env_type] && num_envs > 4096);
}

template <int ENV, int G>
int launch_env_g(urgym_handle h, const KParams& P, cudaStream_t st) {
  static bool attr_done[16] = {false};
  constexpr int EPB = 32 / G;  // envs per 32-thread block
  size_t smem = sizeof(EnvSmem<ENV>) * EPB;
  if (getenv("URGYM_EXTRA_SMEM")) smem += (size_t)atoi(getenv("URGYM_EXTRA_SMEM"));  // occupancy experiments only
  if (!attr_done[h->device & 15]) {
    CK(cudaFuncSetAttribute(env_kernel<ENV, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(env_kernel<ENV, G>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_done[h->device & 15] = true;
  }
  env_kernel<ENV, G><<<(h->N + EPB - 1) / EPB, 32, smem, st>>>(P);
  g_launches++;
  CK(cudaGetLastError());
  return URGYM_OK;
}
template <int ENV>
int launch_env(urgym_handle h, const KParams& P, cudaStream_t st) {
  return launch_env_g<ENV, 32>(h, P, st);
}

// Calls on one handle are ordered even when they arrive on different streams (the *_host entries work on the handle's own
// stream, the device entries on the caller's): the new stream waits for the work the previous call enqueued.
int order_after_last(urgym_handle h, cudaStream_t st) {
  if (h->used && h->last_stream != st) {
    CK(cudaEventRecord(h->order_evt, h->last_stream));
    CK(cudaStreamWaitEvent(st, h->order_evt, 0));
  }
  h->used = true; h->last_stream = st;
  return URGYM_OK;
}

int launch(urgym_handle h, KParams& P, void* stream) {
  CK(cudaSetDevice(h->device));
  { int rc = order_after_last(h, (cudaStream_t)stream); if (rc) return rc; }
  P.model = h->dmodel; P.state = h->state; P.N = h->N; P.contacts = h->contacts; P.ccount = h->ccount; P.hitmask = h->hitmask; P.stats = h->stats;
  P.seed = h->seed; P.gidx0 = h->gidx0; P.sparse = h->sparse;
  cudaStream_t st = (cudaStream_t)stream;  // NULL = the legacy default stream, as in the CUDA runtime
  switch (h->env_type) {
    case 0: return launch_env<0>(h, P, st);
    case 1: return launch_env<1>(h, P, st);
    case 2: return launch_env<2>(h, P, st);
    case 3: return launch_env<3>(h, P, st);
    case 4: return launch_env<4>(h, P, st);
    case 5: return launch_env<5>(h, P, st);
    case 6: return launch_env<6>(h, P, st);
    default: return launch_env<7>(h, P, st);
  }
}

StaticScratch* g_static_ptr[16] = {nullptr};   // per device: the array's address on that device

int static_setup(int device, StaticScratch** out) {
  if (device < 0 || device >= 16) return fail(URGYM_E_ARG, "urgym_step_static: device out of range");
  CK(cudaSetDevice(device));
  static std::mutex mu;                       // two host threads may make their first call on a device at the same time
  std::lock_guard<std::mutex> lock(mu);
  if (!g_static_ptr[device]) {
    StaticScratch* p = nullptr;
    CK(cudaGetSymbolAddress((void**)&p, g_static_scratch));
    for (int t = 0; t < UG_ENV_COUNT; t++) {
      DevModel hm; if (build_model(t, &hm)) return fail(URGYM_E_ARG, "model build failed");
      { const char* e = getenv("URGYM_GENERIC_SWEEPS"); hm.generic_sweeps = e && atoi(e) != 0; }
      CK(cudaMemcpy(&p->model[t], &hm, sizeof(DevModel), cudaMemcpyHostToDevice));
    }
    g_static_ptr[device] = p;
  }
  *out = g_static_ptr[device];
  return URGYM_OK;
}

template <int ENV>
int launch_static(int N, int device, const KParams& P, cudaStream_t st) {
  static bool attr_done[16] = {false};
  const size_t smem = sizeof(EnvSmem<ENV>);
  if (!attr_done[device]) {
    CK(cudaFuncSetAttribute(env_kernel<ENV, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(env_kernel<ENV, 32>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_done[device] = true;
  }
  env_kernel<ENV, 32><<<N, 32, smem, st>>>(P);
  g_launches++;
  CK(cudaGetLastError());
  return URGYM_OK;
}
int launch_static_any(int env_type, int N, int device, const KParams& P, cudaStream_t st) {
  switch (env_type) {
    case 0: return launch_static<0>(N, device, P, st);
    case 1: return launch_static<1>(N, device, P, st);
    case 2: return launch_static<2>(N, device, P, st);
    case 3: return launch_static<3>(N, device, P, st);
    case 4: return launch_static<4>(N, device, P, st);
    case 5: return launch_static<5>(N, device, P, st);
    case 6: return launch_static<6>(N, device, P, st);
    default: return launch_static<7>(N, device, P, st);
  }
}

struct KinCtx { DevModel* dmodel = nullptr; };
KinCtx g_kin[16];
int kin_model(int device, DevModel** out) {
  if (device < 0 || device >= 16) return fail(URGYM_E_ARG, "device out of range");
  CK(cudaSetDevice(device));
  if (!g_kin[device].dmodel) {
    DevModel hm; if (build_model(0, &hm)) return fail(URGYM_E_ARG, "model build failed");
    CK(cudaMalloc(&g_kin[device].dmodel, sizeof(DevModel)));
    CK(cudaMemcpy(g_kin[device].dmodel, &hm, sizeof(DevModel), cudaMemcpyHostToDevice));
  }
  *out = g_kin[device].dmodel;
  return URGYM_OK;
}

}  // namespace

extern "C" {

int urgym_obs_dim(int t) { DevModel m; return build_model(t, &m) ? URGYM_E_ARG : m.obs_dim; }
int urgym_act_dim(int t) { DevModel m; return build_model(t, &m) ? URGYM_E_ARG : m.act_dim; }
int urgym_max_episode_steps(int t) { DevModel m; return build_model(t, &m) ? URGYM_E_ARG : m.max_steps; }
int urgym_num_shapes(int t) { DevModel m; return build_model(t, &m) ? URGYM_E_ARG : m.nshapes; }
int urgym_shape_info(int t, int shape, int* body, int* link) {
  DevModel m; if (build_model(t, &m) || shape < 0 || shape >= m.nshapes) return fail(URGYM_E_ARG, "bad shape index");
  *body = m.shapes[shape].body; *link = m.shapes[shape].link; return URGYM_OK;
}

int urgym_create(int env_type, int num_envs, int device, uint64_t seed, int env_index_offset, int reward_sparse, urgym_handle* out) {
  if (!out || num_envs <= 0 || env_type < 0 || env_type >= UG_ENV_COUNT) return fail(URGYM_E_ARG, "urgym_create: bad arguments");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(URGYM_E_CUDA, "no CUDA device: urgym_b200 has no CPU path");
  if (device < 0 || device >= ndev) return fail(URGYM_E_ARG, "urgym_create: device out of range");
  CK(cudaSetDevice(device));
  if (getenv("URGYM_STACK")) CK(cudaDeviceSetLimit(cudaLimitStackSize, (size_t)atoi(getenv("URGYM_STACK"))));
  urgym_handle h = new urgym_env_s;
  h->env_type = env_type; h->N = num_envs; h->device = device; h->seed = seed; h->gidx0 = env_index_offset; h->sparse = reward_sparse;
  h->split_reset = default_split_reset(env_type, num_envs);
  if (build_model(env_type, &h->hmodel)) { delete h; return fail(URGYM_E_ARG, "model build failed"); }
  { const char* e = getenv("URGYM_GENERIC_SWEEPS"); h->hmodel.generic_sweeps = e && atoi(e) != 0; }
  size_t n = (size_t)num_envs; int od = h->hmodel.obs_dim, ad = h->hmodel.act_dim;
  CK(cudaMalloc(&h->dmodel, sizeof(DevModel)));
  CK(cudaMemcpy(h->dmodel, &h->hmodel, sizeof(DevModel), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&h->state, n * UG_STATE * sizeof(real)));
  CK(cudaMalloc(&h->contacts, n * UG_MAXC * 2 * sizeof(int32_t)));
  CK(cudaMalloc(&h->ccount, n * sizeof(int32_t)));
  CK(cudaMalloc(&h->hitmask, n * UG_HITWORDS * sizeof(uint32_t)));
  CK(cudaMemset(h->hitmask, 0, n * UG_HITWORDS * sizeof(uint32_t)));
  CK(cudaMalloc(&h->stats, 8 * sizeof(double)));
  CK(cudaMemset(h->stats, 0, 8 * sizeof(double)));
  CK(cudaMemset(h->contacts, 0xff, n * UG_MAXC * 2 * sizeof(int32_t)));
  CK(cudaMemset(h->ccount, 0, n * sizeof(int32_t)));
  CK(cudaMalloc(&h->d_action, n * ad * sizeof(real))); CK(cudaMalloc(&h->d_u, n * sizeof(real))); CK(cudaMalloc(&h->d_draws, n * 4 * sizeof(real)));
  CK(cudaMalloc(&h->d_obs, n * od * sizeof(real))); CK(cudaMalloc(&h->d_final, n * od * sizeof(real))); CK(cudaMalloc(&h->d_reward, n * sizeof(real)));
  CK(cudaMalloc(&h->d_term, n)); CK(cudaMalloc(&h->d_trunc, n)); CK(cudaMalloc(&h->d_succ, n)); CK(cudaMalloc(&h->d_mask, n));
  CK(cudaMalloc(&h->d_nticks, n * 2 * sizeof(int32_t)));
  CK(cudaMalloc(&h->d_defer, n)); CK(cudaMemset(h->d_defer, 0, n));
  CK(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&h->order_evt, cudaEventDisableTiming));
  std::vector<real> init(n * UG_STATE, 0._r);
  for (size_t i = 0; i < n; i++) { real* r = init.data() + i * UG_STATE; r[37] = 1._r; for (int b = 0; b < UG_MAXFREE; b++) r[38 + 13 * b + 6] = 1._r; }
  CK(cudaMemcpy(h->state, init.data(), init.size() * sizeof(real), cudaMemcpyHostToDevice));
  *out = h;
  return URGYM_OK;
}

int urgym_destroy(urgym_handle h) {
  if (!h) return URGYM_OK;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  void* ptrs[] = {h->dmodel, h->state, h->contacts, h->ccount, h->hitmask, h->stats, h->d_action, h->d_u, h->d_draws, h->d_obs, h->d_final, h->d_reward, h->d_term, h->d_trunc, h->d_succ, h->d_mask, h->d_nticks, h->d_defer};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  if (h->order_evt) cudaEventDestroy(h->order_evt);
  delete h;
  return URGYM_OK;
}

int urgym_reset(urgym_handle h, const uint8_t* mask_dev, const real* draws_dev, real* obs_dev, int32_t* nticks_dev, void* stream) {
  if (!h || !obs_dev) return fail(URGYM_E_ARG, "urgym_reset: null argument");
  KParams P; std::memset(&P, 0, sizeof(P));
  P.mode = MODE_RESET; P.mask = mask_dev; P.draws = draws_dev; P.obs = obs_dev; P.nticks = nticks_dev;
  return launch(h, P, stream);
}

int urgym_step(urgym_handle h, const real* action_dev, const real* u_dev, const real* draws_dev, real* obs_dev, real* reward_dev,
               uint8_t* terminated_dev, uint8_t* truncated_dev, uint8_t* success_dev, real* final_obs_dev, int32_t* nticks_dev,
               int autoreset, void* stream) {
  if (!h || !action_dev || !obs_dev || !reward_dev || !terminated_dev || !truncated_dev || !success_dev) return fail(URGYM_E_ARG, "urgym_step: null argument");
  KParams P; std::memset(&P, 0, sizeof(P));
  P.mode = MODE_STEP; P.action = action_dev; P.u = u_dev; P.draws = draws_dev; P.obs = obs_dev; P.reward = reward_dev;
  P.term = terminated_dev; P.trunc = truncated_dev; P.succ = success_dev; P.final_obs = final_obs_dev; P.nticks = nticks_dev; P.autoreset = autoreset;
  if (!autoreset || !h->split_reset) return launch(h, P, stream);
  // Two launches, same arithmetic per env: the step grid flags the finished envs, a reset-only grid follows (§4.5).
  P.defer = h->d_defer;
  int rc = launch(h, P, stream);
  if (rc) return rc;
  KParams R; std::memset(&R, 0, sizeof(R));
  R.mode = MODE_RESET; R.mask = h->d_defer; R.draws = draws_dev; R.obs = obs_dev; R.nticks = nticks_dev; R.deferred_pass = 1;
  return launch(h, R, stream);
}

int urgym_step_static(int env_type, int device, int num_envs, uint64_t seed, int env_index_offset, int reward_sparse,
                      const real* state_in_dev, const real* action_dev, const real* u_dev, const real* draws_dev, real* state_out_dev,
                      real* obs_dev, real* reward_dev, uint8_t* terminated_dev, uint8_t* truncated_dev, uint8_t* success_dev,
                      real* final_obs_dev, int32_t* nticks_dev, double* stats_dev, int autoreset, void* stream) {
  if (env_type < 0 || env_type >= UG_ENV_COUNT || num_envs <= 0) return fail(URGYM_E_ARG, "urgym_step_static: bad env type or env count");
  if (num_envs > UG_STATIC_MAX_ENVS)
    return fail(URGYM_E_ARG, "urgym_step_static: num_envs " + std::to_string(num_envs) + " exceeds URGYM_STATIC_MAX_ENVS (" +
                             std::to_string(UG_STATIC_MAX_ENVS) + "), the size of the library's static scratch; split the batch or use urgym_step");
  if (!state_in_dev || !state_out_dev || !action_dev || !obs_dev || !reward_dev || !terminated_dev || !truncated_dev || !success_dev)
    return fail(URGYM_E_ARG, "urgym_step_static: null argument");
  StaticScratch* sc; { int rc = static_setup(device, &sc); if (rc) return rc; }
  cudaStream_t st = (cudaStream_t)stream;
  if (state_out_dev != state_in_dev)
    CK(cudaMemcpyAsync(state_out_dev, state_in_dev, (size_t)num_envs * UG_STATE * sizeof(real), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemsetAsync(sc->stats, 0, sizeof(sc->stats), st));   // the counters accumulate by atomics: they must start at zero
  const bool split = autoreset && default_split_reset(env_type, num_envs);
  KParams P; std::memset(&P, 0, sizeof(P));
  P.model = &sc->model[env_type]; P.state = state_out_dev; P.N = num_envs; P.stats = sc->stats;
  P.seed = seed; P.gidx0 = env_index_offset; P.sparse = reward_sparse;
  P.mode = MODE_STEP; P.action = action_dev; P.u = u_dev; P.draws = draws_dev; P.obs = obs_dev; P.reward = reward_dev;
  P.term = terminated_dev; P.trunc = truncated_dev; P.succ = success_dev; P.final_obs = final_obs_dev; P.nticks = nticks_dev; P.autoreset = autoreset;
  if (split) P.defer = sc->defer;
  { int rc = launch_static_any(env_type, num_envs, device, P, st); if (rc) return rc; }
  if (split) {   // same scheduling as urgym_step: the finished envs restart in a reset-only grid
    KParams R = P;
    R.mode = MODE_RESET; R.mask = sc->defer; R.defer = nullptr; R.deferred_pass = 1;
    R.action = nullptr; R.u = nullptr; R.reward = nullptr; R.term = nullptr; R.trunc = nullptr; R.succ = nullptr; R.final_obs = nullptr; R.autoreset = 0;
    int rc = launch_static_any(env_type, num_envs, device, R, st); if (rc) return rc;
  }
  if (stats_dev) CK(cudaMemcpyAsync(stats_dev, sc->stats, sizeof(sc->stats), cudaMemcpyDeviceToDevice, st));
  return URGYM_OK;
}

int urgym_step_logic(urgym_handle h, const real* action_dev, const real* u_dev, const int32_t* pairs_dev, const int32_t* count_dev,
                     real* obs_dev, real* reward_dev, uint8_t* terminated_dev, uint8_t* truncated_dev, uint8_t* success_dev, void* stream) {
  if (!h || !action_dev || !pairs_dev || !count_dev || !obs_dev || !reward_dev || !terminated_dev || !truncated_dev || !success_dev)
    return fail(URGYM_E_ARG, "urgym_step_logic: null argument");
  KParams P; std::memset(&P, 0, sizeof(P));
  P.mode = MODE_STEP; P.action = action_dev; P.u = u_dev; P.obs = obs_dev; P.reward = reward_dev;
  P.term = terminated_dev; P.trunc = truncated_dev; P.succ = success_dev; P.inject_pairs = pairs_dev; P.inject_count = count_dev;
  return launch(h, P, stream);
}

int urgym_tick(urgym_handle h, int nticks, void* stream) {
  if (!h || nticks <= 0) return fail(URGYM_E_ARG, "urgym_tick: bad arguments");
  KParams P; std::memset(&P, 0, sizeof(P));
  P.mode = MODE_TICKS; P.raw_ticks = nticks;
  return launch(h, P, stream);
}

int urgym_step_host(urgym_handle h, const real* action, const real* u, const real* draws, real* obs, real* reward, uint8_t* terminated,
                    uint8_t* truncated, uint8_t* success, real* final_obs, int32_t* nticks, int autoreset) {
  if (!h || !action || !obs || !reward || !terminated || !truncated || !success) return fail(URGYM_E_ARG, "urgym_step_host: null argument");
  CK(cudaSetDevice(h->device));
  size_t n = (size_t)h->N; int od = h->hmodel.obs_dim, ad = h->hmodel.act_dim; cudaStream_t st = h->own_stream;
  { int rc = order_after_last(h, st); if (rc) return rc; }
  CK(cudaMemcpyAsync(h->d_action, action, n * ad * sizeof(real), cudaMemcpyHostToDevice, st));
  if (u) CK(cudaMemcpyAsync(h->d_u, u, n * sizeof(real), cudaMemcpyHostToDevice, st));
  if (draws) CK(cudaMemcpyAsync(h->d_draws, draws, n * 4 * sizeof(real), cudaMemcpyHostToDevice, st));
  int rc = urgym_step(h, h->d_action, u ? h->d_u : nullptr, draws ? h->d_draws : nullptr, h->d_obs, h->d_reward, h->d_term, h->d_trunc, h->d_succ,
                      h->d_final, h->d_nticks, autoreset, st);
  if (rc) return rc;
  CK(cudaMemcpyAsync(obs, h->d_obs, n * od * sizeof(real), cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(reward, h->d_reward, n * sizeof(real), cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(terminated, h->d_term, n, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(truncated, h->d_trunc, n, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(success, h->d_succ, n, cudaMemcpyDeviceToHost, st));
  if (final_obs) CK(cudaMemcpyAsync(final_obs, h->d_final, n * od * sizeof(real), cudaMemcpyDeviceToHost, st));
  if (nticks) CK(cudaMemcpyAsync(nticks, h->d_nticks, n * 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return URGYM_OK;
}

int urgym_reset_host(urgym_handle h, const uint8_t* mask, const real* draws, real* obs, int32_t* nticks) {
  if (!h || !obs) return fail(URGYM_E_ARG, "urgym_reset_host: null argument");
  CK(cudaSetDevice(h->device));
  size_t n = (size_t)h->N; int od = h->hmodel.obs_dim; cudaStream_t st = h->own_stream;
  { int rc = order_after_last(h, st); if (rc) return rc; }
  if (mask) CK(cudaMemcpyAsync(h->d_mask, mask, n, cudaMemcpyHostToDevice, st));
  if (draws) CK(cudaMemcpyAsync(h->d_draws, draws, n * 4 * sizeof(real), cudaMemcpyHostToDevice, st));
  if (mask) CK(cudaMemcpyAsync(h->d_obs, obs, n * od * sizeof(real), cudaMemcpyHostToDevice, st));  // unmasked rows keep the caller's values
  int rc = urgym_reset(h, mask ? h->d_mask : nullptr, draws ? h->d_draws : nullptr, h->d_obs, h->d_nticks, st);
  if (rc) return rc;
  CK(cudaMemcpyAsync(obs, h->d_obs, n * od * sizeof(real), cudaMemcpyDeviceToHost, st));
  if (nticks) CK(cudaMemcpyAsync(nticks, h->d_nticks, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return URGYM_OK;
}

int urgym_get_state(urgym_handle h, real* state_dev, void* stream) {
  if (!h || !state_dev) return fail(URGYM_E_ARG, "urgym_get_state: null argument");
  CK(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;  // NULL = the legacy default stream, as in the CUDA runtime
  { int rc = order_after_last(h, st); if (rc) return rc; }
  CK(cudaMemcpyAsync(state_dev, h->state, (size_t)h->N * UG_STATE * sizeof(real), cudaMemcpyDeviceToDevice, st));
  return URGYM_OK;
}
int urgym_set_state(urgym_handle h, const real* state_dev, void* stream) {
  if (!h || !state_dev) return fail(URGYM_E_ARG, "urgym_set_state: null argument");
  CK(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;  // NULL = the legacy default stream, as in the CUDA runtime
  { int rc = order_after_last(h, st); if (rc) return rc; }
  CK(cudaMemcpyAsync(h->state, state_dev, (size_t)h->N * UG_STATE * sizeof(real), cudaMemcpyDeviceToDevice, st));
  return URGYM_OK;
}
int urgym_contacts(urgym_handle h, int32_t* pairs_dev, int32_t* count_dev, void* stream) {
  if (!h || !pairs_dev || !count_dev) return fail(URGYM_E_ARG, "urgym_contacts: null argument");
  CK(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;  // NULL = the legacy default stream, as in the CUDA runtime
  { int rc = order_after_last(h, st); if (rc) return rc; }
  CK(cudaMemcpyAsync(pairs_dev, h->contacts, (size_t)h->N * UG_MAXC * 2 * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(count_dev, h->ccount, (size_t)h->N * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  return URGYM_OK;
}

int urgym_contact_pairs(urgym_handle h, uint32_t* mask_dev, void* stream) {
  if (!h || !mask_dev) return fail(URGYM_E_ARG, "urgym_contact_pairs: null argument");
  CK(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  { int rc = order_after_last(h, st); if (rc) return rc; }
  CK(cudaMemcpyAsync(mask_dev, h->hitmask, (size_t)h->N * UG_HITWORDS * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  return URGYM_OK;
}
int urgym_num_pairs(int t) { DevModel m; return build_model(t, &m) ? URGYM_E_ARG : m.npairs; }
int urgym_pair_info(int t, int pair, int* shape_a, int* shape_b) {
  DevModel m; if (build_model(t, &m) || pair < 0 || pair >= m.npairs || !shape_a || !shape_b) return fail(URGYM_E_ARG, "bad pair index");
  *shape_a = m.pairs[pair][0]; *shape_b = m.pairs[pair][1]; return URGYM_OK;
}

int urgym_fk(int device, int n, const real* q_dev, real* pose_dev, void* stream) {
  if (n <= 0 || !q_dev || !pose_dev) return fail(URGYM_E_ARG, "urgym_fk: bad arguments");
  DevModel* dm; int rc = kin_model(device, &dm); if (rc) return rc;
  fk_kernel<<<(n + 63) / 64, 64, 0, (cudaStream_t)stream>>>(dm, n, q_dev, pose_dev); g_launches++;
  CK(cudaGetLastError()); return URGYM_OK;
}
int urgym_ik(int device, int n, const real* q_dev, const real* target_dev, real* out_dev, void* stream) {
  if (n <= 0 || !q_dev || !target_dev || !out_dev) return fail(URGYM_E_ARG, "urgym_ik: bad arguments");
  DevModel* dm; int rc = kin_model(device, &dm); if (rc) return rc;
  ik_kernel<<<(n + 63) / 64, 64, 0, (cudaStream_t)stream>>>(dm, n, q_dev, target_dev, out_dev); g_launches++;
  CK(cudaGetLastError()); return URGYM_OK;
}
int urgym_gripper_center(int device, int n, const real* q_dev, real* out_dev, void* stream) {
  if (n <= 0 || !q_dev || !out_dev) return fail(URGYM_E_ARG, "urgym_gripper_center: bad arguments");
  DevModel* dm; int rc = kin_model(device, &dm); if (rc) return rc;
  CK(cudaFuncSetAttribute(gripper_center_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SmemKin)));
  gripper_center_kernel<<<n, 32, sizeof(SmemKin), (cudaStream_t)stream>>>(dm, n, q_dev, out_dev); g_launches++;
  CK(cudaGetLastError()); return URGYM_OK;
}

int urgym_stats(urgym_handle h, double out[8]) {
  if (!h || !out) return fail(URGYM_E_ARG, "urgym_stats: null argument");
  CK(cudaSetDevice(h->device));
  if (h->last_stream) CK(cudaStreamSynchronize(h->last_stream));
  CK(cudaMemcpy(out, h->stats, 8 * sizeof(double), cudaMemcpyDeviceToHost));
  return URGYM_OK;
}
int urgym_kernel_info(int env_type, int* group_lanes, int* smem_bytes_per_env) {
  if (env_type < 0 || env_type >= UG_ENV_COUNT || !group_lanes || !smem_bytes_per_env) return fail(URGYM_E_ARG, "urgym_kernel_info: bad arguments");
  *group_lanes = group_size();
  const int sz[UG_ENV_COUNT] = {(int)sizeof(EnvSmem<0>), (int)sizeof(EnvSmem<1>), (int)sizeof(EnvSmem<2>), (int)sizeof(EnvSmem<3>), (int)sizeof(EnvSmem<4>),
                                (int)sizeof(EnvSmem<5>), (int)sizeof(EnvSmem<6>), (int)sizeof(EnvSmem<7>)};
  *smem_bytes_per_env = sz[env_type];
  return URGYM_OK;
}
// Development aid (only meaningful in a -DUG_PHASE_TIMING build): summed per-phase warp cycles since the last call.
int urgym_phase_cycles(unsigned long long out[8]) {
#ifdef UG_PHASE_TIMING
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(out, g_phase_cycles, 8 * sizeof(unsigned long long)));
  unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  CK(cudaMemcpyToSymbol(g_phase_cycles, z, sizeof(z)));
  return URGYM_OK;
#else
  for (int i = 0; i < 8; i++) out[i] = 0;
  return URGYM_E_STATE;
#endif
}
long long urgym_launch_count(void) { return g_launches.load(); }
int urgym_real_bytes(void) { return (int)sizeof(real); }
const char* urgym_last_error(void) { return g_err.c_str(); }

}  // extern "C"
