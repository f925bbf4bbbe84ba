// go1sim.cu -- kernels + C ABI (include/go1sim.h) of the batched Go1 simulator for sm_100a.
//
// One control step of every environment is ONE kernel launch (k_step): ONE ENVIRONMENT PER WARP.  The warp
// loads its state row into shared memory, runs frame_skip physics substeps there (go1_warp.cuh), evaluates
// the fused observation / reward / cost / termination logic (go1_mujoco_env.py:483-532), auto-resets a
// finished environment with Philox noise and stores the row back.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo [-DGO1_REAL=double] (see build.py)
#ifdef GO1_EMU
#include "cuda_emu.h"   // tests/emu: CPU SIMT shim used only by the test-suite (never by the product)
#define GO1_DYN_SMEM(name) unsigned char* name = g_emu_block.dyn_smem
#define GO1_LAUNCH(kernel, grid, block, smem, stream, ...) emu_launch(kernel, grid, block, __VA_ARGS__)
#else
#include <cuda_runtime.h>
#define GO1_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#define GO1_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<grid, block, smem, stream>>>(__VA_ARGS__)
#endif
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/go1sim.h"
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
namespace go1 { __device__ unsigned long long g_phase[5][16]; }     // diagnostic build only (tools/phase_times.py)
#endif
#include "go1_warp.cuh"

#ifndef GO1_VARIANT      // timing variants (tools/build_variant.sh) may shrink the capacities
static_assert(go1::CW == GO1_MAX_CONTACTS && go1::PW == GO1_MAX_PAIR_CONTACTS, "include/go1sim.h capacities must match the kernel build");
#endif

#ifndef GO1_REAL
#define GO1_REAL float
#endif
#ifndef GO1_WARPS
#define GO1_WARPS 12         // environments (warps) per block
#endif
#ifndef GO1_MINBLOCKS
#define GO1_MINBLOCKS 2
#endif

namespace go1 {

typedef GO1_REAL real;

// ---------------------------------------------------------------------------------------------
// state slab: one row per environment, reals [N][S_STRIDE] + ints [N][I_STRIDE]
enum {
    S_QPOS = 0, S_QVEL = 19, S_WARM = 37, S_LASTACT = 55, S_AIR = 67, S_DEV = 71, S_VDES = 74, S_EPRET = 77, S_NREAL = 78, S_STRIDE = 80
};
enum { I_STEP = 0, I_UNHEALTHY = 1, I_LASTCON = 2, I_RESETCNT = 3, I_BAD = 4, I_FFT = 5 /* front feet touched (biped) */, I_NINT = 6, I_STRIDE = 8 };
// diag layout (real, per env, [field][N])
enum { D_QACC = 0, D_QFA = 18, D_QFC = 36, D_CFRC = 54, D_TRUNKZ = 138, D_COM = 139, D_NREAL = 142 };
enum { DI_NCON = 0, DI_ITER = 1, DI_ITERTOT = 2, DI_OVERFLOW = 3, DI_NINT = 4 };
// contact capture: per env: count, then CW entries: gid, dim, dist, pos3, force6
enum { DC_STRIDE = 12, DC_ENV = 1 + CW * DC_STRIDE };

template <typename T>
struct KParams {
    const DevModel<T>* model;
    EnvCfg<T> cfg;
    int n_envs;
    T* sreal; int* sint;
    const float* actions;          // [N][12] or null
    int action_mode;               // 0 buffer, 1 random, 2 zero
    const uint8_t* reset_mask;     // reset-only launches
    int reset_only;
    float* obs; float* rew; uint8_t* done; uint8_t* trunc; float* term_obs; float* info;
    T* diag; int* diag_i; T* diag_con;   // nullable
    T* bipk; float* bip_info;      // biped mode only: [N][BK_N] kinematic snapshot, [N][GO1_BIPED_INFO_DIM] reward_info terms
    double* stats;                 // nullable: GO1_NSTATS doubles
    int* queue;                    // environment queue head (zeroed before every launch)
    const int* order;              // queue position -> environment (sorted by last step's solver work)
    int* work;                     // per environment: Newton iterations of this control step
    unsigned long long seed, env_offset, launch_counter;
};

// ---------------------------------------------------------------------------------------------
// Philox4x32-10
struct U4 { unsigned x, y, z, w; };
__device__ __forceinline__ U4 philox(unsigned long long seed, unsigned c0, unsigned c1, unsigned c2, unsigned c3) {
    unsigned k0 = (unsigned)seed, k1 = (unsigned)(seed >> 32);
#pragma unroll 1
    for (int i = 0; i < 10; i++) {
        unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        unsigned n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return U4{c0, c1, c2, c3};
}
__device__ __forceinline__ float u01(unsigned u) { return (float)(u >> 8) * (1.0f / 16777216.0f); }          // [0,1)
__device__ __forceinline__ float u01_open(unsigned u) { return ((float)(u >> 8) + 0.5f) * (1.0f / 16777216.0f); }  // (0,1)

// ---------------------------------------------------------------------------------------------
// env logic helpers (go1_mujoco_env.py)
template <typename T> GO1_DEV void projected_gravity(const T* q, const T* grav, T* pg) {   // :596-608, :1017-1038
    const T w = q[0], x = q[1], y = q[2], z = q[3];
    T t2 = T(2) * (w * y - z * x);
    t2 = t2 > T(1) ? T(1) : t2; t2 = t2 < T(-1) ? T(-1) : t2;
    T e[3];
    if (sizeof(T) == 4) {
        e[0] = atan2f(T(2) * (w * x + y * z), T(1) - T(2) * (x * x + y * y));
        e[1] = asinf(t2);
        e[2] = atan2f(T(2) * (w * z + x * y), T(1) - T(2) * (y * y + z * z));
    } else {
        e[0] = atan2((double)(T(2) * (w * x + y * z)), (double)(T(1) - T(2) * (x * x + y * y)));
        e[1] = asin((double)t2);
        e[2] = atan2((double)(T(2) * (w * z + x * y)), (double)(T(1) - T(2) * (y * y + z * z)));
    }
    const T d = grav[0] * e[0] + grav[1] * e[1] + grav[2] * e[2];
    T v[3] = {d * e[0], d * e[1], d * e[2]};
    const T n = Num<T>::sqrt_(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (n == 0) { pg[0] = v[0]; pg[1] = v[1]; pg[2] = v[2]; }
    else { pg[0] = v[0] / n; pg[1] = v[1] / n; pg[2] = v[2] / n; }
}
template <typename T> GO1_DEV T clipv(T v, T lim) { return v > lim ? lim : (v < -lim ? -lim : v); }

// the 48-d observation (:915-947), two values per lane; `E.last_act` must still hold a_{t-1}
template <typename T>
GO1_DEV void write_obs(const DevModel<T>& m, const EnvCfg<T>& cfg, const WEnv<T>& E, int lane, float* o) {
    T pg[3];
    projected_gravity(E.qpos + 3, m.gravity, pg);
    const T lim = cfg.obs_clip;
#pragma unroll
    for (int rnd = 0; rnd < 2; rnd++) {
        const int i = lane + 32 * rnd;
        if (i >= GO1_OBS_DIM) break;
        T v;
        if (i < 3) v = E.qvel[i] * T(2);
        else if (i < 6) v = E.qvel[i] * T(0.25);
        else if (i < 9) v = i == 6 ? pg[0] : (i == 7 ? pg[1] : pg[2]);
        else if (i < 12) v = E.vdes[i - 9] * T(2);
        else if (i < 24) v = E.qpos[7 + i - 12] - m.key_qpos[7 + i - 12];
        else if (i < 36) v = E.qvel[6 + i - 24] * T(0.05);
        else v = E.last_act[i - 36];
        o[i] = (float)clipv(v, lim);
    }
}

// reset_model (:949-1001) with Philox noise; every lane computes the same draws, lanes split the writes
template <typename T>
GO1_DEV void reset_env(const DevModel<T>& m, const EnvCfg<T>& cfg, WEnv<T>& E, int lane, unsigned long long seed, unsigned long long gid) {
    __syncwarp();
    const int resetcnt = E.ei[I_RESETCNT] + 1;
    __syncwarp();
    const U4 r0 = philox(seed, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)resetcnt, 0x52535430u);
    const U4 r1 = philox(seed, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)resetcnt, 0x52535431u);
    if (lane < 19) {
        T q = m.key_qpos[lane];
        const bool fallen = u01(r0.x) < cfg.fallen_prob;
        if (cfg.biped && !fallen) {      // :967-972: rear-leg stance, trunk pitched -95 deg, dropped from 0.65 m
            if (lane >= 7) q = cfg.biped_ready[lane - 7];
            if (lane == 2) q = cfg.biped_reset_z;
            if (lane == 3) q = cfg.biped_reset_qw;
            if (lane == 5) q = cfg.biped_reset_qy;
            if (lane == 4 || lane == 6) q = 0;
        }
        if (fallen) {
            const float lo = 3.14159265358979f / 2.1f, hi = 3.14159265358979f / 1.5f;
            const float ang = lo + (hi - lo) * u01(r0.y);
            float s, c;
            sincosf(0.5f * ang, &s, &c);
            const bool roll = u01(r0.z) < 0.5f;
            if (lane == 2) q = T(0.1);
            if (lane == 3) q = c;
            if (lane == 4) q = roll ? T(s) : T(0);
            if (lane == 5) q = roll ? T(0) : T(s);
            if (lane == 6) q = 0;
        }
        if (cfg.rand_power > 0 && lane >= 7) {
            const int j = lane - 7, leg = j / 3, i = j - 3 * leg;
            const U4 rn = philox(seed, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)resetcnt, 0x52535440u + leg);
            const float a0 = sqrtf(-2.0f * logf(u01_open(rn.x))), a1 = sqrtf(-2.0f * logf(u01_open(rn.z)));
            float s0, c0, s1, c1;
            sincospif(2.0f * u01(rn.y), &s0, &c0);
            sincospif(2.0f * u01(rn.w), &s1, &c1);
            (void)s1;
            const float nz = i == 0 ? a0 * c0 : (i == 1 ? a0 * s0 : a1 * c1);
            q = q + T(0.1) * cfg.rand_power * T(nz);
            q = q < GO1_LEGF(lim_lo, j) ? GO1_LEGF(lim_lo, j) : q;
            q = q > GO1_LEGF(lim_hi, j) ? GO1_LEGF(lim_hi, j) : q;
        }
        E.qpos[lane] = q;
    }
    if (lane < 18) { E.qvel[lane] = 0; E.warm[lane] = 0; }
    if (lane < 12) E.last_act[lane] = 0;
    if (lane < 4) E.air[lane] = 0;
    if (lane < 3) {
        E.dev[lane] = 0;
        const float uu = lane == 0 ? u01(r0.w) : (lane == 1 ? u01(r1.x) : u01(r1.y));
        E.vdes[lane] = cfg.vdes_min[lane] + (cfg.vdes_max[lane] - cfg.vdes_min[lane]) * T(uu);
    }
    if (lane == 0) {
        E.ep_ret = 0;
        E.ei[I_STEP] = 0; E.ei[I_UNHEALTHY] = 0; E.ei[I_LASTCON] = 0; E.ei[I_RESETCNT] = resetcnt; E.ei[I_FFT] = 0;
    }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(GO1_WARPS * 32, GO1_MINBLOCKS) k_step(const __grid_constant__ KParams<T> P) {
    GO1_DYN_SMEM(smem_raw);
    constexpr size_t MODEL_BYTES = (offsetof(DevModel<T>, gp_a) + 15) & ~size_t(15);      // the geom-pair table stays in global memory
    DevModel<T>& m = *reinterpret_cast<DevModel<T>*>(smem_raw);
    {
        const int* src = reinterpret_cast<const int*>(P.model);
        int* dst = reinterpret_cast<int*>(smem_raw);
        for (int i = threadIdx.x; i < (int)((offsetof(DevModel<T>, gp_a) + 3) / 4); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    WEnv<T>& E = reinterpret_cast<WEnv<T>*>(smem_raw + MODEL_BYTES)[warp];
    const EnvCfg<T>& cfg = P.cfg;
    const int N = P.n_envs;
    typedef Num<T> NN;
    double st_acc = 0;     // lane i < GO1_NSTATS: i-th rollout statistic summed over the environments this warp stepped
    // persistent blocks: the block pulls the next GO1_WARPS environments from a global queue until the slab is done
    __shared__ int s_base;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_base = atomicAdd(P.queue, GO1_WARPS);
        __syncthreads();
        if (s_base >= N) break;
        const int env = s_base + warp < N ? (P.reset_only ? s_base + warp : P.order[s_base + warp]) : N;
        const bool active = env < N && !(P.reset_only && P.reset_mask && !P.reset_mask[env]);
        if (P.reset_only && !active) continue;
        if (!active) {      // keeps the block's barriers balanced
            SubOut<T> dummy;
            dummy.iters_total = 0; dummy.overflow = 0; dummy.bad = 0; dummy.nsub = 0; dummy.qacc = 0; dummy.qacc_prev = 0; dummy.qacc_prev2 = 0;
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
            dummy.tlast = clock64(); dummy.maxmode = 0; dummy.rounds = 0;
            for (int i = 0; i < 7; i++) { dummy.twork[i] = 0; dummy.twait[i] = 0; }
#endif
            for (int s = 0; s < cfg.frame_skip; s++) substep_w(m, cfg, E, lane, dummy, false, (T*)nullptr, false);
            continue;
        }
        const unsigned long long gid = P.env_offset + (unsigned long long)env;
        T* sr = P.sreal + (size_t)env * S_STRIDE;
        int* si = P.sint + (size_t)env * I_STRIDE;
        // ---- load the state row (coalesced)
        T* row = E.qpos;
#pragma unroll 1
        for (int i = lane; i < S_NREAL; i += 32) row[i] = sr[i];
        if (lane < I_NINT) E.ei[lane] = si[lane];
        __syncwarp();
        auto store_state = [&]() {
            __syncwarp();
#pragma unroll 1
            for (int i = lane; i < S_NREAL; i += 32) sr[i] = row[i];
            if (lane < I_NINT) si[lane] = E.ei[lane];
        };
        if (P.reset_only) {
            reset_env(m, cfg, E, lane, P.seed, gid);
            if (lane == 0) E.ei[I_BAD] = 0;
            if (P.obs) write_obs(m, cfg, E, lane, P.obs + (size_t)env * GO1_OBS_DIM);
            store_state();
        } else {
            // ---- action of this control step (lanes 0..11)
            T act = 0;
            if (lane < 12) {
                if (P.action_mode == 0) act = (T)P.actions[(size_t)env * GO1_ACT_DIM + lane];
                else if (P.action_mode == 1) {
                    const int leg = lane / 3, i = lane - 3 * leg;
                    const U4 r = philox(P.seed ^ 0xA5A5A5A55A5A5A5Aull, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)P.launch_counter, 0x41435400u + leg);
                    const float u = i == 0 ? u01(r.x) : (i == 1 ? u01(r.y) : u01(r.z));
                    act = m.ctrl_lo[lane] + (m.ctrl_hi[lane] - m.ctrl_lo[lane]) * T(u);
                }
                E.ctrl[lane] = act;
            }
            // ---- do_simulation: frame_skip x mj_step, then mj_rnePostConstraint (:481)
            SubOut<T> out;
            out.iters_total = 0; out.overflow = 0; out.bad = 0; out.ncon = 0; out.iters = 0;
            out.qacc = 0; out.qfa = 0; out.qfc = 0; out.trunk_z = 0; out.qacc_prev = 0; out.qacc_prev2 = 0; out.nsub = 0;
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
            out.tlast = clock64(); out.maxmode = 0; out.rounds = 0;
            for (int i = 0; i < 7; i++) { out.twork[i] = 0; out.twait[i] = 0; }
#endif
            const bool capture = P.diag_con != nullptr;
            T* con_out = capture ? P.diag_con + (size_t)env * DC_ENV : (T*)nullptr;
            {
                // mj_checkPos / mj_checkVel: a non-finite or exploded state is never simulated
                int pre = 0;
                for (int i = lane; i < S_LASTACT; i += 32) pre |= !(NN::abs_(row[i]) <= T(1e10));
                if (__any_sync(FULL, pre)) out.bad = 1;
            }
            if (lane == 0) E.ei[I_STEP] += 1;
            __syncwarp();
            if (out.bad) {   // mjData of a skipped step: nothing the reward code may read is left uninitialised
                for (int i = lane; i < NB * 6; i += 32) (&E.cfrc[0][0])[i] = 0;
                if (lane < 3) E.com[lane] = 0;
                if (lane == 0) { E.ncon = 0; E.nfl = 0; E.overflow = 0; }
                __syncwarp();
            }
            for (int s = 0; s < cfg.frame_skip; s++) {
                const bool last = s == cfg.frame_skip - 1;
                substep_w(m, cfg, E, lane, out, capture && last, con_out, !out.bad,
                          (cfg.biped && last) ? P.bipk + (size_t)env * BK_N : (T*)nullptr);
            }
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
            if (lane == 0) {
                const int cls = (E.ncon > E.nfl && out.maxmode == 0) ? 1 : (out.maxmode == 0 ? 0 : out.maxmode + 1);     // 0 no pair rows, 1 inactive pair rows, 2 arrow, 3 pair-leg, 4 dense
                for (int i = 0; i < 7; i++) { atomicAdd(&g_phase[cls][i], (unsigned long long)out.twork[i]); atomicAdd(&g_phase[cls][8 + i], (unsigned long long)out.twait[i]); }
                atomicAdd(&g_phase[cls][7], 1ull); atomicAdd(&g_phase[cls][15], (unsigned long long)out.rounds);
            }
#endif
            if (out.bad) { out.qacc = 0; out.qfa = 0; out.qfc = 0; out.trunk_z = 0; }
            const bool legdof = lane >= 6 && lane < NDOF;
            const int jl = legdof ? lane - 6 : 0;
            // ---- health (:433-471) on the post-step state; non-finite states also raise the sticky bad flag
            int nonfinite = 0;
            for (int i = lane; i < S_WARM; i += 32) nonfinite |= !(NN::abs_(row[i]) <= T(1e10));
            nonfinite = __any_sync(FULL, nonfinite) | out.bad;
            const T qz = E.qpos[2], qroll = E.qpos[4], qpitch = E.qpos[5];
            int reason = 0;
            if (nonfinite) reason = 1;
            else if (!(cfg.healthy_z[0] <= qz && qz <= cfg.healthy_z[1])) reason = 2;
            else if (!(cfg.healthy_roll[0] <= qroll && qroll <= cfg.healthy_roll[1])) reason = 3;
            else if (!(cfg.healthy_pitch[0] <= qpitch && qpitch <= cfg.healthy_pitch[1])) reason = 4;
            const bool ok = reason == 0;
            T pg[3];
            projected_gravity(E.qpos + 3, m.gravity, pg);
            // ---- rewards (:791-913)
            const T* rw = cfg.reward_w; const T* cw = cfg.cost_w;
            const T ex = E.vdes[0] - E.qvel[0], ey = E.vdes[1] - E.qvel[1], ez = E.vdes[2] - E.qvel[5];
            T lin_r, ang_r;
            if (sizeof(T) == 4) { lin_r = expf(-(ex * ex + ey * ey) / cfg.sigma) * rw[0]; ang_r = expf(-(ez * ez) / cfg.sigma) * rw[1]; }
            else { lin_r = exp((double)(-(ex * ex + ey * ey) / cfg.sigma)) * rw[0]; ang_r = exp((double)(-(ez * ez) / cfg.sigma)) * rw[1]; }
            const T healthy_r = ok ? rw[2] : T(0);
            // feet air time (:633-668): lanes 0..3 = feet (calf bodies 3, 6, 9, 12)
            T air_l = 0, fmag = 0;
            int curr = 0;
            const int lastcon = E.ei[I_LASTCON];
            if (lane < 4) {
                const T* w = E.cfrc[3 + 3 * lane];
                fmag = NN::sqrt_(dot6(w, w));
                curr = fmag > T(1);
            }
            if (lane < 4 && !cfg.biped) {
                const int filt = curr | ((lastcon >> lane) & 1);
                T air = E.air[lane];
                const T first = (air > 0 && filt) ? T(1) : T(0);
                air += cfg.dt;
                const T ts = air - T(0.2) > 0 ? air - T(0.2) : T(0);
                air_l = ts * ts * first;
                if (filt) air = 0;
                E.air[lane] = air;
            }
            const int newcon = (int)(__ballot_sync(FULL, curr) & 0xFu);
            T air_r = warp_sum(air_l);
            if (!(NN::sqrt_(E.vdes[0] * E.vdes[0] + E.vdes[1] * E.vdes[1]) > T(0.1))) air_r = 0;
            if (cfg.biped) air_r = (((newcon >> 2) ^ (newcon >> 3)) & 1) ? T(1) : T(0);     // :646-649: rear feet alternate; no bookkeeping
            air_r *= rw[3];
            // recovery (:534-580)
            T dev[3] = {0, 0, 0};
            if (!(cfg.healthy_z[0] <= qz && qz <= cfg.healthy_z[1])) dev[0] = NN::min_(NN::abs_(qz - cfg.healthy_z[0]), NN::abs_(qz - cfg.healthy_z[1]));
            if (!(cfg.healthy_roll[0] <= qroll && qroll <= cfg.healthy_roll[1])) dev[1] = NN::min_(NN::abs_(qroll - cfg.healthy_roll[0]), NN::abs_(qroll - cfg.healthy_roll[1]));
            if (!(cfg.healthy_pitch[0] <= qpitch && qpitch <= cfg.healthy_pitch[1])) dev[2] = NN::min_(NN::abs_(qpitch - cfg.healthy_pitch[0]), NN::abs_(qpitch - cfg.healthy_pitch[1]));
            T rec = 0;
            if (!ok) rec = (E.dev[0] - dev[0]) + (E.dev[1] - dev[1]) + (E.dev[2] - dev[2]);
            const T rec_r = rec * rw[4];
            T gu = 0;
            if (!ok) {
                T hr = out.trunk_z / cfg.healthy_z[0];
                hr = hr < 0 ? T(0) : (hr > 1 ? T(1) : hr);
                gu = hr * T(1.5) + (T(1) - (pg[0] * pg[0] + pg[1] * pg[1])) * T(0.5);
            }
            const T gu_r = gu * rw[5];
            T rewards = lin_r + ang_r + healthy_r + air_r + rec_r + gu_r;
            // costs: per-joint terms on the dof lanes, reduced across the warp
            T s_tau = 0, s_rate = 0, s_jl = 0, s_jv = 0, s_acc = 0, s_djp = 0, s_coll = 0;
            const T act_j = __shfl_sync(FULL, act, jl);      // action of this lane's joint
            if (legdof) {
                const T ql = E.qpos[7 + jl], vl = E.qvel[6 + jl];
                s_tau = out.qfa * out.qfa;
                const T drr = E.last_act[jl] - act_j;
                s_rate = drr * drr;
                const T lo = m.soft_lo[jl] - ql, hi = ql - m.soft_hi[jl];
                s_jl = (lo > 0 ? lo : T(0)) + (hi > 0 ? hi : T(0));
                s_jv = vl * vl;
                s_acc = out.qacc * out.qacc / (NN::abs_(vl) + T(1e-6));
                const T dj = ql - m.key_ctrl[jl];
                s_djp = dj * dj;
            }
            for (int i = lane; i < 48; i += 32) {     // hips + thighs (:124-126, :688-693)
                const int k = i / 12, rem = i - 12 * k;
                const T v = E.cfrc[1 + 3 * k + rem / 6][rem % 6];
                s_coll += v * v;
            }
            s_tau = warp_sum(s_tau); s_rate = warp_sum(s_rate); s_jl = warp_sum(s_jl); s_jv = warp_sum(s_jv);
            s_acc = warp_sum(s_acc); s_djp = warp_sum(s_djp); s_coll = warp_sum(s_coll);
            const T disc = ok ? T(1) : T(0.1);
            const T ctrl_c = s_tau * disc * cw[0];
            const T vert_c = E.qvel[2] * E.qvel[2] * cw[1];
            const T xyang_c = (E.qvel[3] * E.qvel[3] + E.qvel[4] * E.qvel[4]) * cw[2];
            const T rate_c = s_rate * disc * cw[3];
            const T jl_c = s_jl * cw[4];
            const T jv_c = s_jv * cw[5];
            const T acc_c = s_acc * disc * cw[6];
            const T orient_c = (pg[0] * pg[0] + pg[1] * pg[1]) * cw[7];
            const T coll_c = (NN::sqrt_(s_coll) > T(0.1) ? T(1) : T(0)) * cw[8];
            const T djp_c = s_djp * cw[9];
            const T unh_c = (ok ? T(0) : T(1)) * cw[10];
            T costs = ctrl_c + rate_c + vert_c + xyang_c + jl_c + jv_c + acc_c + coll_c + unh_c;
            // ---- bipedal mode (:869-904): one reward and nine costs replace orientation / default_joint_position
            T bterm[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
            if (cfg.biped) {
                const T* bw = cfg.biped_w;
                __syncwarp();
                T bk[BK_N];
                const T* bsrc = P.bipk + (size_t)env * BK_N;
#pragma unroll
                for (int i = 0; i < BK_N; i++) bk[i] = out.bad ? T(0) : bsrc[i];
                const T s_fc = warp_sum(lane < 2 ? fmag * fmag : T(0));                                       // :426-431
                const bool rear_air = (__ballot_sync(FULL, lane >= 2 && lane < 4 && fmag < T(1)) & 0xCu) == 0xCu;   // :880-882
                T s_uc = 0, s_ab = 0;
                for (int i = lane; i < 42; i += 32) {       // trunk, the four thighs, the front calves (:192-200, :772-781)
                    const int k = i / 6;
                    const int b = k == 0 ? 0 : (k <= 4 ? 2 + 3 * (k - 1) : 3 + 3 * (k - 5));
                    const T v = E.cfrc[b][i - 6 * k];
                    s_uc += v * v;
                }
                if (lane < 4) { const T qa = E.qpos[7 + 3 * lane]; s_ab = qa * qa; }                           // :759-769
                s_uc = warp_sum(s_uc); s_ab = warp_sum(s_ab);
                auto pos = [](T v) { return v > 0 ? v : T(0); };
                bterm[0] = pos(bk[BK_TRUNK_XZ]) * bw[0];                                                        // :388-401
                bterm[1] = s_fc * bw[1];
                bterm[2] = rear_air ? bw[2] : T(0);
                bterm[3] = (pos(bk[BK_FRONT_CALF_Z] - bk[BK_FRONT_FOOT_Z]) + pos(bk[BK_FRONT_CALF_Z + 1] - bk[BK_FRONT_FOOT_Z + 1])) * bw[3];   // :403-423
                bterm[4] = pos(bk[BK_REAR_HIP_Y] - bk[BK_REAR_HIP_Y + 1]) * bw[4];                              // :315-331
                bterm[5] = (pos(T(0.2) - bk[BK_HIP_Z + 2]) + pos(T(0.2) - bk[BK_HIP_Z + 3])) * T(10) * bw[5];   // :334-351
                const T d0 = pos(bk[BK_HIP_Z] - bk[BK_FRONT_FOOT_Z]), d1 = pos(bk[BK_HIP_Z + 1] - bk[BK_FRONT_FOOT_Z + 1]);
                bterm[6] = (d0 * d0 + d1 * d1) * bw[6];                                                         // :354-376
                bterm[7] = s_ab * bw[7];
                bterm[8] = s_uc * bw[8];
                // :269-312 self_collision_cost: contacts between the front-right and front-left limbs, or rear-right and rear-left
                // (hip, thigh, calf of legs FR 0 / FL 1 / RR 2 / RL 3), counted over data.contact of the last substep
                int nself = 0;
                if (!out.bad)
                    for (int c = E.nfl + lane; c < E.ncon; c += 32) {
                        const int mt = E.c_meta[c];
                        const int b2 = meta_b2(mt), b1 = meta_b1p(mt) - 1;
                        if (b1 >= 1 && b2 >= 1) {
                            const int l1 = (b1 - 1) / 3, l2 = (b2 - 1) / 3;
                            nself += ((l1 ^ l2) == 1 && (l1 >> 1) == (l2 >> 1)) ? 1 : 0;
                        }
                    }
                nself = warp_sum(nself);
                bterm[9] = T(nself) * bw[9];
                rewards += bterm[0];
#pragma unroll
                for (int i = 1; i < 10; i++) costs += bterm[i];
            } else {
                costs = costs + orient_c + djp_c;
            }
            T reward = rewards - costs;
            reward = reward > 0 ? reward : T(0);
            if (nonfinite) reward = 0;
            // ---- termination (:487-498)
            const int unhealthy = ok ? 0 : E.ei[I_UNHEALTHY] + 1;
            const int ep_len = E.ei[I_STEP];
            const bool terminated = unhealthy > cfg.max_unhealthy_steps;
            const bool truncated = ep_len >= cfg.max_episode_steps;
            const bool done = terminated || truncated || nonfinite;
            const T ep_ret = E.ep_ret + reward;
            const float x_pos = (float)E.qpos[0], y_pos = (float)E.qpos[1];
            __syncwarp();
            if (lane == 0) {
                E.ei[I_UNHEALTHY] = unhealthy; E.ep_ret = ep_ret;
                if (!cfg.biped) E.ei[I_LASTCON] = newcon;
                if (ok) { E.dev[0] = E.dev[1] = E.dev[2] = 0; } else { E.dev[0] = dev[0]; E.dev[1] = dev[1]; E.dev[2] = dev[2]; }
            }
            // ---- optional capture of mjData (tests)
            if (P.diag) {
                T* dg = P.diag + env;
                if (lane < NDOF) {
                    dg[(size_t)(D_QACC + lane) * N] = out.qacc; dg[(size_t)(D_QFA + lane) * N] = out.qfa; dg[(size_t)(D_QFC + lane) * N] = out.qfc;
                }
                for (int i = lane; i < 6; i += 32) dg[(size_t)(D_CFRC + i) * N] = 0;   // world body
                for (int i = lane; i < NB * 6; i += 32) dg[(size_t)(D_CFRC + 6 + i) * N] = (&E.cfrc[0][0])[i];
                if (lane < 3) dg[(size_t)(D_COM + lane) * N] = out.bad ? T(0) : E.com[lane] + E.qpos[lane];
                if (lane == 0) {
                    dg[(size_t)D_TRUNKZ * N] = out.trunk_z;
                    int* di = P.diag_i + env;
                    di[(size_t)DI_NCON * N] = out.ncon; di[(size_t)DI_ITER * N] = out.iters; di[(size_t)DI_ITERTOT * N] = out.iters_total;
                    di[(size_t)DI_OVERFLOW * N] = out.overflow;
                }
            }
            // ---- outputs + auto-reset (SubprocVecEnv worker semantics); obs uses a_{t-1} (:483 runs before :529)
            float* obs_row = P.obs + (size_t)env * GO1_OBS_DIM;
            float* tobs_row = P.term_obs ? P.term_obs + (size_t)env * GO1_OBS_DIM : nullptr;
            if (done) { if (tobs_row) write_obs(m, cfg, E, lane, tobs_row); }
            else write_obs(m, cfg, E, lane, obs_row);
            __syncwarp();
            if (lane < 12) E.last_act[lane] = act;    // :529
            if (lane == 0) {
                P.rew[env] = (float)reward;
                // queue key of the next control step (k_order: heaviest first).  Environments with robot-robot contacts go to the front
                // in one group: the pair phases of a Newton round (pair Hessian, coupled-leg factorisation) then run in blocks whose
                // warps mostly all need them, and the other blocks never wait for them.
                {
#if !defined(GO1_WORK_KEY) || GO1_WORK_KEY == 3
                    // inside a level: by the number of robot-robot contacts, then floor contacts (the per-phase cost of a round follows
                    // the row counts; measured 2 % better than the iteration-based key, profiles/r2_experiments.md section 3)
                    int w = 8 * (E.ncon - E.nfl) + E.nfl;
#elif GO1_WORK_KEY == 0      // timing experiments only: other orders inside a level
                    int w = (out.iters_total + 2 * out.ncon) >> 1;
#elif GO1_WORK_KEY == 1
                    int w = 4 * out.ncon + (out.iters_total >> 4);
#elif GO1_WORK_KEY == 2
                    int w = out.iters_total;
#endif
                    w = w > 63 ? 63 : w;
#ifndef GO1_NO_PAIR_SORT
                    // level 0: floor contacts only; 1: robot-robot contacts inside one leg / with the trunk; 2: between ONE pair of legs
                    // (6x6 path); 3: between several pairs (dense path)
                    int lvl = 0;
                    int code0 = -1;
                    if (!out.bad)
                        for (int c = E.nfl; c < E.ncon; c++) {
                            const int mt = E.c_meta[c];
                            const int b2 = meta_b2(mt), b1 = meta_b1p(mt) - 1;
                            const bool cross = b1 >= 1 && b2 >= 1 && (b1 - 1) / 3 != (b2 - 1) / 3;
                            if (cross) {
                                const int l1 = (b1 - 1) / 3, l2 = (b2 - 1) / 3, code = l1 < l2 ? 4 * l1 + l2 : 4 * l2 + l1;
                                if (code0 < 0) code0 = code;
                                lvl = code != code0 ? 3 : (lvl > 2 ? lvl : 2);
                            } else lvl = lvl > 1 ? lvl : 1;
                        }
                    w += 64 * lvl;
#endif
                    P.work[env] = w;
                }
                P.done[env] = done ? 1 : 0;
                P.trunc[env] = (truncated && !terminated && !nonfinite) ? 1 : 0;
                if (P.info) {
                    float* inf = P.info + (size_t)env * GO1_INFO_DIM;
                    inf[0] = x_pos; inf[1] = y_pos; inf[2] = (float)lin_r; inf[3] = (float)(-ctrl_c); inf[4] = (float)healthy_r;
                    inf[5] = (float)rec_r; inf[6] = (float)gu_r; inf[7] = (float)(-unh_c); inf[8] = (float)(-orient_c); inf[9] = (float)(-djp_c);
                    inf[10] = (float)ep_ret; inf[11] = (float)ep_len; inf[12] = (float)reason; inf[13] = ok ? 1.f : 0.f;
                    inf[14] = (float)out.iters_total; inf[15] = (float)out.ncon;
                    // the value _get_health_status reports in termination_details (:444-459), robot-robot contacts and dropped contacts
                    inf[16] = reason == 2 ? (float)qz : (reason == 3 ? (float)qroll : (reason == 4 ? (float)qpitch : 0.f));
                    inf[17] = out.bad ? 0.f : (float)(E.ncon - E.nfl); inf[18] = (float)out.overflow; inf[19] = 0.f;
                    if (cfg.biped) inf[8] = inf[9] = 0.f;      // the two terms biped mode drops from reward_info (:905-909)
                }
                if (cfg.biped) {
                    const int touched = E.ei[I_FFT];            // latched from the forces of EARLIER steps (:477-479 runs before do_simulation)
                    float* bi = P.bip_info + (size_t)env * GO1_BIPED_INFO_DIM;
                    bi[0] = (float)bterm[0];
#pragma unroll
                    for (int i = 1; i < 10; i++) bi[i] = (float)(-bterm[i]);
                    bi[10] = (truncated && !touched) ? 1.f : 0.f;        // bipedal_success (:520-521)
                    const int now = touched | ((newcon & 3) ? 1 : 0);
                    bi[11] = (float)now;
                    E.ei[I_FFT] = now;
                }
            }
            {
                double v = 0;
                if (lane == 0) v = (double)reward;
                else if (lane == 1) v = done ? 1.0 : 0.0;
                else if (lane == 2) v = done ? (double)ep_len : 0.0;
                else if (lane == 3) v = done ? (double)ep_ret : 0.0;
                else if (lane == 4) v = terminated ? 1.0 : 0.0;
                else if (lane == 5) v = (truncated && !terminated) ? 1.0 : 0.0;
                else if (lane == 6) v = nonfinite ? 1.0 : 0.0;
                else if (lane == 7) v = 1.0;
                else if (lane == 8) v = (double)out.overflow;                 // contacts dropped for lack of slots (summed over substeps)
                else if (lane == 9) v = (double)out.ncon;                      // active contacts, last substep
                else if (lane == 10) v = (double)out.iters_total;              // Newton iterations of the control step
                else if (lane == 11) v = out.bad ? 0.0 : (double)(E.ncon - E.nfl);   // robot-robot contacts, last substep
                st_acc += v;
            }
            if (done) {
                reset_env(m, cfg, E, lane, P.seed, gid);
                if (lane == 0 && nonfinite) E.ei[I_BAD] += 1;
                write_obs(m, cfg, E, lane, obs_row);
            }
            store_state();
        }
    }
    if (P.stats && lane < GO1_NSTATS && st_acc != 0.0) atomicAdd(P.stats + lane, st_acc);
}

// Queue order for the next control step: environments sorted by the solver work of the last one (counting sort,
// heaviest first).  The warps of a block advance in lockstep, so a block is as slow as its slowest environment:
// neighbours in the queue should need about the same number of Newton iterations.  Results do not depend on the order.
__global__ void k_order(const int* __restrict__ work, int* __restrict__ order, int n) {
#ifdef GO1_DBG_IDENTITY_ORDER      // timing experiment only: queue in slab order
    for (int i = threadIdx.x; i < n; i += blockDim.x) order[i] = i;
    return;
#endif
    __shared__ int hist[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) { int k = work[i]; k = k < 0 ? 0 : (k > 255 ? 255 : k); atomicAdd(&hist[255 - k], 1); }
    __syncthreads();
    if (threadIdx.x == 0) { int acc = 0; for (int b = 0; b < 256; b++) { const int c = hist[b]; hist[b] = acc; acc += c; } }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) { int k = work[i]; k = k < 0 ? 0 : (k > 255 ? 255 : k); order[atomicAdd(&hist[255 - k], 1)] = i; }
}

// Spreads the episode phases of a slab that was reset all at once (bench / throughput runs): an environment that is currently
// unhealthy gets its consecutive-unhealthy-step counter (go1_mujoco_env.py:487-493) set to a Philox draw in [0, max_unhealthy],
// so that terminations arrive at a uniform rate instead of all at control step ~800.  Nothing else of the state changes.
__global__ void k_stagger(int* __restrict__ sint, int n, int max_unhealthy, unsigned long long seed, unsigned long long env_offset) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    int* si = sint + (size_t)e * I_STRIDE;
    if (si[I_UNHEALTHY] <= 0) return;
    const unsigned long long gid = env_offset + (unsigned long long)e;
    const U4 r = philox(seed ^ 0x5354414747455221ull, (unsigned)gid, (unsigned)(gid >> 32), 0u, 0x53544752u);
    const int k = (int)(r.x % (unsigned)(max_unhealthy + 1));
    si[I_UNHEALTHY] = k;
    if (si[I_STEP] < k) si[I_STEP] = k;
}

// float32 -> float64 widening of the observation batch (go1_step_host_f64): SB3's observation dtype is float64
// (gymnasium Box of the reference env), and 0.8 M conversions cost 0.5 ms on one host core but nothing here.
__global__ void k_widen(const float* __restrict__ src, double* __restrict__ dst, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = (double)src[i];
}

}  // namespace go1
#include "go1_policy_tc.cuh"
#ifndef GO1_EMU
#include "go1_render.cuh"
#endif
namespace go1 {

// 48-64-64-12 tanh MLP + Gaussian sampling + clip: one thread per env (weights in shared memory).
// Scalar reference of the tensor-core kernel (go1_policy_tc.cuh): used by the CPU emulation build and by go1_policy_act_scalar (tests).
// Round-1 version on the FP32 pipe; the tensor-core (tcgen05) version is tracked in DESIGN.md "next".
__global__ void __launch_bounds__(128) k_policy(const float* __restrict__ obs, float* __restrict__ act, int n, Go1Policy pol,
                                                 const float* lo, const float* hi, unsigned long long seed, unsigned long long counter,
                                                 unsigned long long env_offset) {
    __shared__ float w0[64 * 48], b0[64], w1[64 * 64], b1[64], w2[12 * 64], b2[12], ls[12], slo[12], shi[12];
    for (int i = threadIdx.x; i < 64 * 48; i += blockDim.x) w0[i] = pol.w0[i];
    for (int i = threadIdx.x; i < 64 * 64; i += blockDim.x) w1[i] = pol.w1[i];
    for (int i = threadIdx.x; i < 12 * 64; i += blockDim.x) w2[i] = pol.w2[i];
    for (int i = threadIdx.x; i < 64; i += blockDim.x) { b0[i] = pol.b0[i]; b1[i] = pol.b1[i]; }
    for (int i = threadIdx.x; i < 12; i += blockDim.x) { b2[i] = pol.b2[i]; ls[i] = pol.log_std ? pol.log_std[i] : -1e30f; slo[i] = lo[i]; shi[i] = hi[i]; }
    __syncthreads();
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    float x[48], h0[64], h1[64];
    for (int i = 0; i < 48; i++) x[i] = obs[(size_t)e * 48 + i];
    for (int o = 0; o < 64; o++) { float s = b0[o]; for (int i = 0; i < 48; i++) s += w0[o * 48 + i] * x[i]; h0[o] = tanhf(s); }
    for (int o = 0; o < 64; o++) { float s = b1[o]; for (int i = 0; i < 64; i++) s += w1[o * 64 + i] * h0[i]; h1[o] = tanhf(s); }
    const unsigned long long gid = env_offset + e;
    for (int o = 0; o < 12; o += 4) {
        const U4 r = philox(seed ^ 0x504F4C4943590000ull, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)counter, 0x50000000u + o);
        const float a0 = sqrtf(-2.0f * logf(u01_open(r.x))), a1 = sqrtf(-2.0f * logf(u01_open(r.z)));
        float s0, c0, s1, c1;
        sincospif(2.0f * u01(r.y), &s0, &c0);
        sincospif(2.0f * u01(r.w), &s1, &c1);
        const float nz[4] = {a0 * c0, a0 * s0, a1 * c1, a1 * s1};
        for (int q = 0; q < 4; q++) {
            float s = b2[o + q];
            for (int i = 0; i < 64; i++) s += w2[(o + q) * 64 + i] * h1[i];
            if (pol.log_std) s += expf(ls[o + q]) * nz[q];
            s = fminf(shi[o + q], fmaxf(slo[o + q], s));
            act[(size_t)e * 12 + o + q] = s;
        }
    }
}

}  // namespace go1


// =============================================================================================
// host side / C ABI
using namespace go1;

struct Go1Sim {
    int n_envs = 0, device = 0;
    EnvCfg<real> cfg;
    DevModel<double> hmodel;
    DevModel<real>* dmodel = nullptr;
    DevModel<float>* dmodel_f32 = nullptr;      // renderer's view of the model (== dmodel in the fp32 build)
    int* d_envids = nullptr; unsigned char* d_rgb = nullptr; size_t rgb_cap = 0; int envids_cap = 0;
    real* sreal = nullptr; int* sint = nullptr;
    real* diag = nullptr; int* diag_i = nullptr; real* diag_con = nullptr;
    real* bipk = nullptr; float* bip_info = nullptr; real* bip_ready = nullptr;      // biped mode only
    bool diag_enabled = false;
    double* stats = nullptr;
    int* queue = nullptr;
    int *order = nullptr, *work = nullptr;
    int grid_blocks = 0;
    float *d_act = nullptr, *d_obs = nullptr, *d_rew = nullptr, *d_tobs = nullptr, *d_info = nullptr;
    uint8_t *d_done = nullptr, *d_trunc = nullptr, *d_mask = nullptr;
    float *h_act = nullptr, *h_obs = nullptr, *h_rew = nullptr, *h_tobs = nullptr, *h_info = nullptr;   // pinned staging
    double *d_obs64 = nullptr, *h_obs64 = nullptr;      // go1_step_host_f64 only (allocated on first use)
    uint8_t *h_done = nullptr, *h_trunc = nullptr, *h_mask = nullptr;
    float *d_lo = nullptr, *d_hi = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    unsigned long long seed = 0, env_offset = 0, launch_counter = 0;
    float last_ms = 0; int last_launches = 0;
    std::vector<std::pair<char*, size_t>> pinned;     // caller buffers page-locked through go1_host_register
    std::string err;
};

static std::string g_create_err;

static size_t step_smem_bytes() {
    return ((offsetof(DevModel<real>, gp_a) + 15) & ~size_t(15)) + sizeof(WEnv<real>) * GO1_WARPS;
}

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { h->err = std::string(#call) + ": " + cudaGetErrorString(e_); return GO1_ECUDA; } } while (0)

extern "C" {

void go1_default_config(Go1EnvConfig* c) {
    std::memset(c, 0, sizeof(*c));
    c->biped = 0; c->frame_skip = 10; c->max_episode_steps = 6000; c->max_unhealthy_steps = 750;
    c->rand_power = 0.0f; c->fallen_start_prob = 0.2f;
    c->desired_vel_min[0] = -0.5f; c->desired_vel_max[0] = 0.5f;
    const float rw[6] = {2.0f, 1.0f, 1.0f, 5.0f, 10.0f, 20.0f};
    const float cw[11] = {0.0002f, 2.0f, 0.05f, 0.01f, 10.0f, 0.01f, 2.0e-4f, 1.0f, 1.0f, 0.1f, 5.0f};
    std::memcpy(c->reward_w, rw, sizeof(rw)); std::memcpy(c->cost_w, cw, sizeof(cw));
    c->healthy_z[0] = 0.22f; c->healthy_z[1] = 1.8f;
    c->healthy_roll[0] = -1.3962634015954636f; c->healthy_roll[1] = 1.3962634015954636f;
    c->healthy_pitch[0] = -3.14159265358979323846f; c->healthy_pitch[1] = 0.0f;
    c->tracking_sigma = 0.25f; c->obs_clip = 100.0f; c->solver_tolerance = -1.0f; c->solver_iterations = 0;
    const float bw[10] = {15.0f, 50.0f, 5.0f, 8.0f, 5.0f, 9.0f, 6.0f, 0.7f, 150.0f, 25.0f};
    const float ready[12] = {0.0f, 4.0f, -2.0f, 0.0f, 4.0f, -2.0f, 0.0f, 2.8f, -1.2f, 0.0f, 2.8f, -1.2f};
    std::memcpy(c->biped_w, bw, sizeof(bw)); std::memcpy(c->biped_ready_joints, ready, sizeof(ready));
    c->biped_reset_z = 0.65f; c->biped_reset_pitch_deg = -95.0f;
}

int go1_precision_bits(void) { return (int)sizeof(real) * 8; }
const char* go1_version(void) { return "go1sim 0.1 (sm_100a)"; }
const char* go1_last_error(go1_handle h) { return h ? h->err.c_str() : g_create_err.c_str(); }

static void fill_cfg(EnvCfg<real>& k, const Go1EnvConfig& c, const DevModel<double>& M) {
    k.frame_skip = c.frame_skip; k.max_episode_steps = c.max_episode_steps; k.max_unhealthy_steps = c.max_unhealthy_steps;
    k.solver_iterations = c.solver_iterations > 0 ? c.solver_iterations : M.iterations;
    k.dt = (real)(M.timestep * (double)c.frame_skip);
    k.rand_power = c.rand_power; k.fallen_prob = c.fallen_start_prob;
    for (int i = 0; i < 3; i++) { k.vdes_min[i] = c.desired_vel_min[i]; k.vdes_max[i] = c.desired_vel_max[i]; }
    // weights / ranges are the reference's fp64 literals: widen from the float config, but keep exact
    // double values for the defaults (0.0002f etc. are not representable)
    Go1EnvConfig def; go1_default_config(&def);
    const double rwd[6] = {2.0, 1.0, 1.0, 5.0, 10.0, 20.0};
    const double cwd[11] = {0.0002, 2.0, 0.05, 0.01, 10.0, 0.01, 2.0e-4, 1.0, 1.0, 0.1, 5.0};
    for (int i = 0; i < 6; i++) k.reward_w[i] = (real)(c.reward_w[i] == def.reward_w[i] ? rwd[i] : (double)c.reward_w[i]);
    for (int i = 0; i < 11; i++) k.cost_w[i] = (real)(c.cost_w[i] == def.cost_w[i] ? cwd[i] : (double)c.cost_w[i]);
    const double hz[2] = {0.22, 1.8}, hr[2] = {-1.3962634015954636, 1.3962634015954636}, hp[2] = {-3.14159265358979323846, 0.0};
    for (int i = 0; i < 2; i++) {
        k.healthy_z[i] = (real)(c.healthy_z[i] == def.healthy_z[i] ? hz[i] : (double)c.healthy_z[i]);
        k.healthy_roll[i] = (real)(c.healthy_roll[i] == def.healthy_roll[i] ? hr[i] : (double)c.healthy_roll[i]);
        k.healthy_pitch[i] = (real)(c.healthy_pitch[i] == def.healthy_pitch[i] ? hp[i] : (double)c.healthy_pitch[i]);
    }
    k.sigma = (real)c.tracking_sigma; k.obs_clip = (real)c.obs_clip;
    k.tolerance = (real)(c.solver_tolerance >= 0 ? (double)c.solver_tolerance : M.tolerance);
    k.biped = c.biped ? 1 : 0;
    const double bwd[10] = {15.0, 50.0, 5.0, 8.0, 5.0, 9.0, 6.0, 0.7, 150.0, 25.0};
    for (int i = 0; i < 10; i++) k.biped_w[i] = (real)(c.biped_w[i] == def.biped_w[i] ? bwd[i] : (double)c.biped_w[i]);
    k.biped_ready = nullptr;   // device copy made by go1_create
    k.gp_tab = nullptr;        // points into the device model (go1_create)
    k.biped_reset_z = (real)(c.biped_reset_z == def.biped_reset_z ? 0.65 : (double)c.biped_reset_z);
    const double half = 0.5 * (double)c.biped_reset_pitch_deg * (3.14159265358979323846 / 180.0);     // np.deg2rad(-95) / 2 (:970-971)
    k.biped_reset_qw = (real)std::cos(half); k.biped_reset_qy = (real)std::sin(half);
    for (int f = 0; f < 2; f++) for (int e = 0; e < 3; e++) k.foot_site[f][e] = 0;
}

int go1_create(const void* model_blob, size_t nbytes, const Go1EnvConfig* cfg_in, int n_envs, int cuda_device, uint64_t seed,
               uint64_t env_id_offset, go1_handle* out) {
    if (!out || !model_blob || n_envs <= 0) { g_create_err = "go1_create: bad argument"; return GO1_EINVAL; }
    *out = nullptr;
    Go1EnvConfig cfg;
    if (cfg_in) cfg = *cfg_in; else go1_default_config(&cfg);
    if (cfg.frame_skip <= 0 || cfg.max_episode_steps <= 0) { g_create_err = "bad config"; return GO1_EINVAL; }
    Go1Sim* h = new Go1Sim();
    if (!build_dev_model(model_blob, nbytes, h->hmodel, h->err)) { g_create_err = "model rejected: " + h->err; delete h; return GO1_EMODEL; }
    h->n_envs = n_envs; h->device = cuda_device; h->seed = seed; h->env_offset = env_id_offset;
    fill_cfg(h->cfg, cfg, h->hmodel);
    if (cfg.biped) {
        double fs[NLEG][3];
        if (!read_foot_sites(model_blob, nbytes, fs, h->err)) { g_create_err = "model rejected: " + h->err; delete h; return GO1_EMODEL; }
        for (int f = 0; f < 2; f++) for (int e = 0; e < 3; e++) h->cfg.foot_site[f][e] = (real)fs[f][e];
    }
    auto fail = [&](const char* what, cudaError_t e) { g_create_err = std::string(what) + ": " + cudaGetErrorString(e); go1_destroy(h); return GO1_ECUDA; };
    cudaError_t e;
    if ((e = cudaSetDevice(cuda_device)) != cudaSuccess) return fail("cudaSetDevice", e);
    if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    cudaEventCreate(&h->ev0); cudaEventCreate(&h->ev1);
    DevModel<real>* hm = new DevModel<real>();
    cast_dev_model(h->hmodel, *hm);
    if ((e = cudaMalloc(&h->dmodel, sizeof(DevModel<real>))) != cudaSuccess) { delete hm; return fail("cudaMalloc model", e); }
    e = cudaMemcpy(h->dmodel, hm, sizeof(DevModel<real>), cudaMemcpyHostToDevice);
    delete hm;
    if (e != cudaSuccess) return fail("cudaMemcpy model", e);
    h->cfg.gp_tab = reinterpret_cast<const unsigned char*>(h->dmodel) + offsetof(DevModel<real>, gp_a);
    const size_t N = (size_t)n_envs;
#define ALLOC(p, bytes) if ((e = cudaMalloc((void**)&(p), (bytes))) != cudaSuccess) return fail("cudaMalloc " #p, e); cudaMemset((p), 0, (bytes))
    ALLOC(h->sreal, sizeof(real) * S_STRIDE * N); ALLOC(h->sint, sizeof(int) * I_STRIDE * N);
    ALLOC(h->stats, sizeof(double) * GO1_NSTATS); ALLOC(h->queue, sizeof(int) * 4); ALLOC(h->order, sizeof(int) * N); ALLOC(h->work, sizeof(int) * N);
    ALLOC(h->d_act, sizeof(float) * 12 * N); ALLOC(h->d_obs, sizeof(float) * 48 * N); ALLOC(h->d_rew, sizeof(float) * N);
    ALLOC(h->d_tobs, sizeof(float) * 48 * N); ALLOC(h->d_info, sizeof(float) * GO1_INFO_DIM * N);
    ALLOC(h->d_done, N); ALLOC(h->d_trunc, N); ALLOC(h->d_mask, N);
    ALLOC(h->d_lo, sizeof(float) * 12); ALLOC(h->d_hi, sizeof(float) * 12);
    if (cfg.biped) {
        ALLOC(h->bipk, sizeof(real) * BK_N * N); ALLOC(h->bip_info, sizeof(float) * GO1_BIPED_INFO_DIM * N); ALLOC(h->bip_ready, sizeof(real) * 12);
        real rdy[12];
        for (int i = 0; i < 12; i++) rdy[i] = (real)(double)cfg.biped_ready_joints[i];     // 4.0, 2.8f ...: widened from float, |error| <= 1e-7 rad
        const double exact[12] = {0.0, 4.0, -2.0, 0.0, 4.0, -2.0, 0.0, 2.8, -1.2, 0.0, 2.8, -1.2};
        Go1EnvConfig def; go1_default_config(&def);
        for (int i = 0; i < 12; i++) if (cfg.biped_ready_joints[i] == def.biped_ready_joints[i]) rdy[i] = (real)exact[i];
        if ((e = cudaMemcpy(h->bip_ready, rdy, sizeof(rdy), cudaMemcpyHostToDevice)) != cudaSuccess) return fail("cudaMemcpy biped_ready", e);
        h->cfg.biped_ready = h->bip_ready;
    }
#undef ALLOC
#define HALLOC(p, bytes) if ((e = cudaMallocHost((void**)&(p), (bytes))) != cudaSuccess) return fail("cudaMallocHost " #p, e)
    HALLOC(h->h_act, sizeof(float) * 12 * N); HALLOC(h->h_obs, sizeof(float) * 48 * N); HALLOC(h->h_rew, sizeof(float) * N);
    HALLOC(h->h_tobs, sizeof(float) * 48 * N); HALLOC(h->h_info, sizeof(float) * GO1_INFO_DIM * N);
    HALLOC(h->h_done, N); HALLOC(h->h_trunc, N); HALLOC(h->h_mask, N);
#undef HALLOC
    float lo[12], hi[12];
    for (int i = 0; i < 12; i++) { lo[i] = (float)h->hmodel.ctrl_lo[i]; hi[i] = (float)h->hmodel.ctrl_hi[i]; }
    cudaMemcpy(h->d_lo, lo, sizeof(lo), cudaMemcpyHostToDevice); cudaMemcpy(h->d_hi, hi, sizeof(hi), cudaMemcpyHostToDevice);
    if ((e = cudaFuncSetAttribute(k_step<real>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)step_smem_bytes())) != cudaSuccess)
        return fail("cudaFuncSetAttribute", e);
#ifdef GO1_EMU
    h->grid_blocks = 64;
#else
    {
        int per_sm = 0, sms = 0;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_step<real>, GO1_WARPS * 32, step_smem_bytes())) != cudaSuccess) return fail("occupancy query", e);
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cuda_device)) != cudaSuccess) return fail("device attribute", e);
        if (per_sm < 1) { g_create_err = "k_step does not fit on an SM"; go1_destroy(h); return GO1_ECUDA; }
        h->grid_blocks = per_sm * sms;      // persistent warps: one resident wave
    }
#endif
    *out = h;
    int rc = go1_reset(h, nullptr, nullptr, nullptr);
    if (rc != GO1_OK) { g_create_err = h->err; go1_destroy(h); *out = nullptr; return rc; }
    if ((e = cudaStreamSynchronize(h->stream)) != cudaSuccess) { *out = nullptr; return fail("initial reset", e); }
    return GO1_OK;
}

int go1_destroy(go1_handle h) {
    if (!h) return GO1_EINVAL;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    void* dev[] = {h->order, h->work, h->queue, h->dmodel, h->sreal, h->sint, h->diag, h->diag_i, h->diag_con, h->bipk, h->bip_info, h->bip_ready, h->stats, h->d_act, h->d_obs, h->d_rew, h->d_tobs,
                   h->d_info, h->d_done, h->d_trunc, h->d_mask, h->d_lo, h->d_hi};
    for (void* p : dev) if (p) cudaFree(p);
    if (h->dmodel_f32 && (void*)h->dmodel_f32 != (void*)h->dmodel) cudaFree(h->dmodel_f32);
    if (h->d_envids) cudaFree(h->d_envids);
    if (h->d_rgb) cudaFree(h->d_rgb);
    if (h->d_obs64) cudaFree(h->d_obs64);
    if (h->h_obs64) cudaFreeHost(h->h_obs64);
    for (auto& r : h->pinned) cudaHostUnregister(r.first);
    h->pinned.clear();
    void* host[] = {h->h_act, h->h_obs, h->h_rew, h->h_tobs, h->h_info, h->h_done, h->h_trunc, h->h_mask};
    for (void* p : host) if (p) cudaFreeHost(p);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return GO1_OK;
}

int go1_num_envs(go1_handle h) { return h ? h->n_envs : GO1_EINVAL; }

int go1_action_bounds(go1_handle h, float* low12, float* high12) {
    if (!h || !low12 || !high12) return GO1_EINVAL;
    for (int i = 0; i < 12; i++) { low12[i] = (float)h->hmodel.ctrl_lo[i]; high12[i] = (float)h->hmodel.ctrl_hi[i]; }
    return GO1_OK;
}

static KParams<real> base_params(go1_handle h) {
    KParams<real> P;
    std::memset(&P, 0, sizeof(P));
    P.model = h->dmodel; P.cfg = h->cfg; P.n_envs = h->n_envs; P.sreal = h->sreal; P.sint = h->sint;
    P.seed = h->seed; P.env_offset = h->env_offset; P.launch_counter = h->launch_counter; P.queue = h->queue; P.order = h->order; P.work = h->work;
    if (h->diag_enabled) { P.diag = h->diag; P.diag_i = h->diag_i; P.diag_con = h->diag_con; }
    P.bipk = h->bipk; P.bip_info = h->bip_info;
    return P;
}
static int launch(go1_handle h, const KParams<real>& P, cudaStream_t s) {
    int grid = (h->n_envs + GO1_WARPS - 1) / GO1_WARPS;
    if (grid > h->grid_blocks) grid = h->grid_blocks;
    CK(cudaMemsetAsync(h->queue, 0, sizeof(int), s));
    if (!P.reset_only) {
#ifdef GO1_EMU
        GO1_LAUNCH(k_order, 1, 32, 0, s, h->work, h->order, h->n_envs);
#else
        GO1_LAUNCH(k_order, 1, 1024, 0, s, h->work, h->order, h->n_envs);
#endif
    }
    GO1_LAUNCH(k_step<real>, grid, GO1_WARPS * 32, step_smem_bytes(), s, P);
    CK(cudaGetLastError());
    return GO1_OK;
}

int go1_reset(go1_handle h, const uint8_t* mask_dev, float* obs_dev, void* stream) {
    if (!h) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    KParams<real> P = base_params(h);
    P.reset_only = 1; P.reset_mask = mask_dev; P.obs = obs_dev;
    return launch(h, P, stream ? (cudaStream_t)stream : h->stream);
}

int go1_step(go1_handle h, const float* actions_dev, float* obs_dev, float* rew_dev, uint8_t* done_dev, uint8_t* trunc_dev,
             float* terminal_obs_dev, float* info_dev, void* stream) {
    if (!h || !actions_dev || !obs_dev || !rew_dev || !done_dev || !trunc_dev) { if (h) h->err = "go1_step: null buffer"; return GO1_EINVAL; }
    CK(cudaSetDevice(h->device));
    KParams<real> P = base_params(h);
    P.actions = actions_dev; P.action_mode = 0; P.obs = obs_dev; P.rew = rew_dev; P.done = done_dev; P.trunc = trunc_dev;
    P.term_obs = terminal_obs_dev; P.info = info_dev;
    h->launch_counter++;
    return launch(h, P, stream ? (cudaStream_t)stream : h->stream);
}

// ---------------------------------------------------------------------------------------------------------------------
// go1_step_noalloc: go1_step for callers that must not reach a device allocator (CUDA graph capture, tight serving loops).
// The step kernel's scratch -- the claim counter of the persistent work queue, the queue order k_order writes and the per-environment
// work keys k_step leaves for the next order -- lives in ONE file-scope device array instead of the handle's cudaMalloc'ed
// buffers.  Every handle on a device shares it: calls that overlap on the GPU (two streams) race on it.  The array keeps the last
// call's contents: the counter is cleared by every call; the work keys are deliberately NOT (they are only a sorting hint; k_order
// clamps whatever it finds and emits a valid permutation, and the results of a step do not depend on the queue order).
// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address: the address the kernels get comes
// from cudaGetSymbolAddress, once per device, with that device current.
constexpr int NOALLOC_HDR = 64;      // ints before the two arrays (the queue counter is word 0): keeps the arrays 256-byte aligned
#ifdef GO1_EMU
alignas(256) static int g_noalloc_scratch[NOALLOC_HDR + 2 * GO1_NOALLOC_MAX_ENVS];      // CPU emulation build (tests only): plain host memory
#else
static __device__ __align__(256) int g_noalloc_scratch[NOALLOC_HDR + 2 * GO1_NOALLOC_MAX_ENVS];      // 128.25 KB per GPU that loads the library
#endif
static std::mutex g_noalloc_mu;
static int* g_noalloc_dev[64] = {};  // device address of g_noalloc_scratch, per device ordinal

static int noalloc_scratch(go1_handle h, int** out) {
    if (h->device < 0 || h->device >= 64) { h->err = "go1_step_noalloc: device ordinal beyond 63"; return GO1_EINVAL; }
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_dev[h->device]) {
#ifdef GO1_EMU
        g_noalloc_dev[h->device] = g_noalloc_scratch;      // CPU emulation build (tests only): device memory is host memory
#else
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);      // current device = h->device (set by the caller below)
        if (e != cudaSuccess || !ptr) {
            h->err = std::string("go1_step_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e);
            return GO1_ECUDA;
        }
        g_noalloc_dev[h->device] = (int*)ptr;
#endif
    }
    *out = g_noalloc_dev[h->device];
    return GO1_OK;
}

int go1_step_noalloc(go1_handle h, const float* actions_dev, float* obs_dev, float* rew_dev, uint8_t* done_dev, uint8_t* trunc_dev,
                     float* terminal_obs_dev, float* info_dev, void* stream) {
    if (!h || !actions_dev || !obs_dev || !rew_dev || !done_dev || !trunc_dev) { if (h) h->err = "go1_step_noalloc: null buffer"; return GO1_EINVAL; }
    if (h->n_envs > GO1_NOALLOC_MAX_ENVS) {
        h->err = "go1_step_noalloc: the handle has " + std::to_string(h->n_envs) + " environments, the static scratch holds GO1_NOALLOC_MAX_ENVS = " +
                 std::to_string(GO1_NOALLOC_MAX_ENVS) + " (use go1_step, or several handles stepped one after the other)";
        return GO1_EINVAL;
    }
    CK(cudaSetDevice(h->device));
    int* sc = nullptr;
    const int rc = noalloc_scratch(h, &sc);
    if (rc != GO1_OK) return rc;
    KParams<real> P = base_params(h);
    P.queue = sc; P.order = sc + NOALLOC_HDR; P.work = sc + NOALLOC_HDR + GO1_NOALLOC_MAX_ENVS;
    P.actions = actions_dev; P.action_mode = 0; P.obs = obs_dev; P.rew = rew_dev; P.done = done_dev; P.trunc = trunc_dev;
    P.term_obs = terminal_obs_dev; P.info = info_dev;
    h->launch_counter++;
    cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
    int grid = (h->n_envs + GO1_WARPS - 1) / GO1_WARPS;
    if (grid > h->grid_blocks) grid = h->grid_blocks;
    CK(cudaMemsetAsync(P.queue, 0, sizeof(int), s));        // the claim counter must start at zero; nothing else has to
#ifdef GO1_EMU
    GO1_LAUNCH(k_order, 1, 32, 0, s, (const int*)P.work, sc + NOALLOC_HDR, h->n_envs);
#else
    GO1_LAUNCH(k_order, 1, 1024, 0, s, (const int*)P.work, sc + NOALLOC_HDR, h->n_envs);
#endif
    GO1_LAUNCH(k_step<real>, grid, GO1_WARPS * 32, step_smem_bytes(), s, P);
    CK(cudaGetLastError());
    return GO1_OK;
}

// true when [p, p + bytes) lies inside a buffer the caller page-locked with go1_host_register: DMA goes straight to it
static bool caller_pinned(go1_handle h, const void* p, size_t bytes) {
    const char* c = (const char*)p;
    for (const auto& r : h->pinned)
        if (c >= r.first && c + bytes <= r.first + r.second) return true;
    return false;
}

int go1_host_register(go1_handle h, void* ptr, size_t bytes) {
    if (!h || !ptr || !bytes) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    CK(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
    h->pinned.emplace_back((char*)ptr, bytes);
    return GO1_OK;
}

int go1_host_unregister(go1_handle h, void* ptr) {
    if (!h || !ptr) return GO1_EINVAL;
    for (size_t i = 0; i < h->pinned.size(); i++)
        if (h->pinned[i].first == (char*)ptr) {
            CK(cudaSetDevice(h->device));
            if (h->stream) CK(cudaStreamSynchronize(h->stream));
            h->pinned.erase(h->pinned.begin() + i);
            CK(cudaHostUnregister(ptr));
            return GO1_OK;
        }
    h->err = "go1_host_unregister: pointer was not registered";
    return GO1_EINVAL;
}

static int step_host_impl(go1_handle h, const float* actions, float* obs, double* obs64, float* rew, uint8_t* done, uint8_t* trunc, float* terminal_obs, float* info) {
    if (!h || !actions || (!obs && !obs64) || !rew || !done || !trunc) { if (h) h->err = "go1_step_host: null buffer"; return GO1_EINVAL; }
    CK(cudaSetDevice(h->device));
    const size_t N = (size_t)h->n_envs;
    const size_t b_act = sizeof(float) * 12 * N, b_obs = sizeof(float) * 48 * N, b_rew = sizeof(float) * N, b_info = sizeof(float) * GO1_INFO_DIM * N;
    // pageable caller buffers are staged through the library's pinned ones; registered buffers are the DMA endpoints themselves
    const float* src_act = actions;
    if (!caller_pinned(h, actions, b_act)) { std::memcpy(h->h_act, actions, b_act); src_act = h->h_act; }
    float* dst_obs = obs ? (caller_pinned(h, obs, b_obs) ? obs : h->h_obs) : nullptr;
    double* dst_obs64 = nullptr;
    if (obs64) {
        if (!h->d_obs64) {
            if (cudaMalloc((void**)&h->d_obs64, 2 * b_obs) != cudaSuccess || cudaMallocHost((void**)&h->h_obs64, 2 * b_obs) != cudaSuccess) {
                h->err = "go1_step_host_f64: out of memory"; return GO1_ENOMEM;
            }
        }
        dst_obs64 = caller_pinned(h, obs64, 2 * b_obs) ? obs64 : h->h_obs64;
    }
    float* dst_rew = caller_pinned(h, rew, b_rew) ? rew : h->h_rew;
    uint8_t* dst_done = caller_pinned(h, done, N) ? done : h->h_done;
    uint8_t* dst_trunc = caller_pinned(h, trunc, N) ? trunc : h->h_trunc;
    float* dst_tobs = terminal_obs ? (caller_pinned(h, terminal_obs, b_obs) ? terminal_obs : h->h_tobs) : nullptr;
    float* dst_info = info ? (caller_pinned(h, info, b_info) ? info : h->h_info) : nullptr;
    CK(cudaMemcpyAsync(h->d_act, src_act, b_act, cudaMemcpyHostToDevice, h->stream));
    int rc = go1_step(h, h->d_act, h->d_obs, h->d_rew, h->d_done, h->d_trunc, h->d_tobs, info ? h->d_info : nullptr, h->stream);
    if (rc != GO1_OK) return rc;
    if (dst_obs) CK(cudaMemcpyAsync(dst_obs, h->d_obs, b_obs, cudaMemcpyDeviceToHost, h->stream));
    if (dst_obs64) {
        const int nel = (int)(48 * N);
#ifdef GO1_EMU
        GO1_LAUNCH(k_widen, nel, 1, 0, h->stream, h->d_obs, h->d_obs64, nel);
#else
        GO1_LAUNCH(k_widen, (nel + 255) / 256, 256, 0, h->stream, h->d_obs, h->d_obs64, nel);
#endif
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(dst_obs64, h->d_obs64, 2 * b_obs, cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaMemcpyAsync(dst_rew, h->d_rew, b_rew, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(dst_done, h->d_done, N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(dst_trunc, h->d_trunc, N, cudaMemcpyDeviceToHost, h->stream));
    if (dst_tobs) CK(cudaMemcpyAsync(dst_tobs, h->d_tobs, b_obs, cudaMemcpyDeviceToHost, h->stream));
    if (dst_info) CK(cudaMemcpyAsync(dst_info, h->d_info, b_info, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (obs && dst_obs != obs) std::memcpy(obs, dst_obs, b_obs);
    if (obs64 && dst_obs64 != obs64) std::memcpy(obs64, dst_obs64, 2 * b_obs);
    if (dst_rew != rew) std::memcpy(rew, dst_rew, b_rew);
    if (dst_done != done) std::memcpy(done, dst_done, N);
    if (dst_trunc != trunc) std::memcpy(trunc, dst_trunc, N);
    if (terminal_obs && dst_tobs != terminal_obs) std::memcpy(terminal_obs, dst_tobs, b_obs);
    if (info && dst_info != info) std::memcpy(info, dst_info, b_info);
    return GO1_OK;
}

int go1_step_host(go1_handle h, const float* actions, float* obs, float* rew, uint8_t* done, uint8_t* trunc, float* terminal_obs, float* info) {
    if (!obs) { if (h) h->err = "go1_step_host: null buffer"; return GO1_EINVAL; }
    return step_host_impl(h, actions, obs, nullptr, rew, done, trunc, terminal_obs, info);
}

int go1_step_host_f64(go1_handle h, const float* actions, double* obs64, float* rew, uint8_t* done, uint8_t* trunc, float* terminal_obs, float* info) {
    if (!obs64) { if (h) h->err = "go1_step_host_f64: null buffer"; return GO1_EINVAL; }
    return step_host_impl(h, actions, nullptr, obs64, rew, done, trunc, terminal_obs, info);
}

int go1_reset_host(go1_handle h, const uint8_t* mask, float* obs) {
    if (!h) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    const size_t N = (size_t)h->n_envs;
    const uint8_t* dmask = nullptr;
    if (mask) {
        std::memcpy(h->h_mask, mask, N);
        CK(cudaMemcpyAsync(h->d_mask, h->h_mask, N, cudaMemcpyHostToDevice, h->stream));
        dmask = h->d_mask;
    }
    int rc = go1_reset(h, dmask, obs ? h->d_obs : nullptr, h->stream);
    if (rc != GO1_OK) return rc;
    if (obs) CK(cudaMemcpyAsync(h->h_obs, h->d_obs, sizeof(float) * 48 * N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (obs) {
        if (!mask) std::memcpy(obs, h->h_obs, sizeof(float) * 48 * N);
        else for (size_t i = 0; i < N; i++) if (mask[i]) std::memcpy(obs + 48 * i, h->h_obs + 48 * i, sizeof(float) * 48);
    }
    return GO1_OK;
}

static int launch_policy(go1_handle h, const Go1Policy* policy, const Go1Critic* critic, const float* obs_dev, float* act_dev,
                         float* values_dev, float* logp_dev, float* raw_dev, int n, cudaStream_t s);

int go1_rollout(go1_handle h, int n_steps, int action_mode, const Go1Policy* policy, Go1Stats* out, void* stream) {
    if (!h || n_steps < 0) return GO1_EINVAL;
    if (action_mode == GO1_ACTION_POLICY && !policy) { h->err = "go1_rollout: policy mode needs a policy"; return GO1_EINVAL; }
    CK(cudaSetDevice(h->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
    CK(cudaMemsetAsync(h->stats, 0, sizeof(double) * GO1_NSTATS, s));
    int launches = 0;
    CK(cudaEventRecord(h->ev0, s));
    for (int t = 0; t < n_steps; t++) {
        KParams<real> P = base_params(h);
        P.obs = h->d_obs; P.rew = h->d_rew; P.done = h->d_done; P.trunc = h->d_trunc; P.term_obs = nullptr; P.info = nullptr;
        P.stats = h->stats;
        if (action_mode == GO1_ACTION_POLICY) {
            int prc = launch_policy(h, policy, nullptr, h->d_obs, h->d_act, nullptr, nullptr, nullptr, h->n_envs, s);
            if (prc != GO1_OK) return prc;
            launches++;
            P.actions = h->d_act; P.action_mode = 0;
        } else {
            P.action_mode = action_mode == GO1_ACTION_RANDOM ? 1 : 2;
        }
        h->launch_counter++;
        int rc = launch(h, P, s);
        if (rc != GO1_OK) return rc;
        launches += 2;      // k_order + k_step
    }
    CK(cudaEventRecord(h->ev1, s));
    CK(cudaStreamSynchronize(s));
    CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
    h->last_launches = launches;
    if (out) {
        double st[GO1_NSTATS];
        CK(cudaMemcpy(st, h->stats, sizeof(st), cudaMemcpyDeviceToHost));
        out->sum_reward = st[0]; out->n_episodes = st[1]; out->sum_episode_length = st[2]; out->sum_episode_return = st[3];
        out->n_terminated = st[4]; out->n_truncated = st[5]; out->n_bad_state = st[6]; out->env_steps = st[7];
        out->n_contact_overflow = st[8]; out->sum_contacts = st[9]; out->sum_solver_iterations = st[10]; out->sum_self_contacts = st[11];
    }
    return GO1_OK;
}

int go1_stagger_phases(go1_handle h) {
    if (!h) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
#ifdef GO1_EMU
    GO1_LAUNCH(k_stagger, h->n_envs, 1, 0, h->stream, h->sint, h->n_envs, h->cfg.max_unhealthy_steps, h->seed, h->env_offset);
#else
    GO1_LAUNCH(k_stagger, (h->n_envs + 255) / 256, 256, 0, h->stream, h->sint, h->n_envs, h->cfg.max_unhealthy_steps, h->seed, h->env_offset);
#endif
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(h->stream));
    return GO1_OK;
}

// actor (+ critic) forward on n observation rows: tcgen05 kernel on the GPU, the scalar kernel in the CPU emulation build (tests only)
static int launch_policy(go1_handle h, const Go1Policy* policy, const Go1Critic* critic, const float* obs_dev, float* act_dev,
                         float* values_dev, float* logp_dev, float* raw_dev, int n, cudaStream_t s) {
#ifdef GO1_EMU
    (void)critic; (void)values_dev; (void)logp_dev; (void)raw_dev;
    GO1_LAUNCH(k_policy, n, 1, 0, s, obs_dev, act_dev, n, *policy, h->d_lo, h->d_hi, h->seed, h->launch_counter, h->env_offset);
#else
    static int sm_count[64] = {};       // per device ordinal; 0 = this device has not been set up yet
    const int dv = h->device >= 0 && h->device < 64 ? h->device : 0;
    if (!sm_count[dv]) {                // function attributes are per device: once on each device that launches the kernel
        CK(cudaFuncSetAttribute(k_policy_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::SM_TOTAL));
        CK(cudaDeviceGetAttribute(&sm_count[dv], cudaDevAttrMultiProcessorCount, h->device));
    }
    PolicyTCParams P;
    P.obs = obs_dev; P.act = act_dev; P.value = critic ? values_dev : nullptr; P.logp = logp_dev; P.raw_act = raw_dev; P.n = n;
    P.aw0 = policy->w0; P.ab0 = policy->b0; P.aw1 = policy->w1; P.ab1 = policy->b1; P.aw2 = policy->w2; P.ab2 = policy->b2; P.log_std = policy->log_std;
    P.cw0 = critic ? critic->w0 : nullptr; P.cb0 = critic ? critic->b0 : nullptr; P.cw1 = critic ? critic->w1 : nullptr;
    P.cb1 = critic ? critic->b1 : nullptr; P.cw2 = critic ? critic->w2 : nullptr; P.cb2 = critic ? critic->b2 : nullptr;
    P.lo = h->d_lo; P.hi = h->d_hi; P.seed = h->seed; P.counter = h->launch_counter; P.env_offset = h->env_offset;
    // persistent CTAs (one per SM, 1 resident by shared memory): two groups of four warps, each walking its own tiles of 128 rows
    const int tiles = (n + 127) / 128;
    if (tiles > 0) k_policy_tc<<<tiles < sm_count[dv] ? tiles : sm_count[dv], tc::THREADS, tc::SM_TOTAL, s>>>(P);
    CK(cudaGetLastError());
#endif
    return GO1_OK;
}

int go1_policy_forward(go1_handle h, const Go1Policy* actor, const Go1Critic* critic, const float* obs_dev, float* act_dev,
                       float* values_dev, float* log_probs_dev, float* raw_act_dev, int n, void* stream) {
    if (!h || !actor || !obs_dev || !act_dev || n < 0) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    if (n == 0) return GO1_OK;
#ifdef GO1_EMU
    if (critic || values_dev || log_probs_dev || raw_act_dev) { h->err = "go1_policy_forward: values / log-probs / raw actions need the tensor-core kernel (not in the CPU emulation build)"; return GO1_EUNSUPPORTED; }
#endif
    int rc = launch_policy(h, actor, critic, obs_dev, act_dev, values_dev, log_probs_dev, raw_act_dev, n, stream ? (cudaStream_t)stream : h->stream);
    h->launch_counter++;
    return rc;
}

// the round-1 scalar kernel, kept as the fp32 reference the tensor-core kernel is tested against on the GPU
int go1_policy_act_scalar(go1_handle h, const Go1Policy* policy, const float* obs_dev, float* act_dev, int n, void* stream) {
    if (!h || !policy || !obs_dev || !act_dev || n < 0) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
    if (n == 0) return GO1_OK;
#ifdef GO1_EMU
    const int tb = 1;
#else
    const int tb = 128;
#endif
    GO1_LAUNCH(k_policy, (n + tb - 1) / tb, tb, 0, s, obs_dev, act_dev, n, *policy, h->d_lo, h->d_hi, h->seed, h->launch_counter, h->env_offset);
    h->launch_counter++;
    return GO1_OK;
}

int go1_policy_act(go1_handle h, const Go1Policy* policy, const float* obs_dev, float* act_dev, int n, void* stream) {
    if (!h || !policy || !obs_dev || !act_dev || n < 0) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
    if (n == 0) return GO1_OK;
    int rc = launch_policy(h, policy, nullptr, obs_dev, act_dev, nullptr, nullptr, nullptr, n, s);
    h->launch_counter++;
    return rc;
}

int go1_last_rollout_timing(go1_handle h, float* ms, int* launches) {
    if (!h) return GO1_EINVAL;
    if (ms) *ms = h->last_ms;
    if (launches) *launches = h->last_launches;
    return GO1_OK;
}

// state access: one contiguous row per environment (test/diagnostic path, synchronous)
static int copy_row(go1_handle h, bool to_device, void* base, size_t row_bytes, size_t bytes, int env, void* host) {
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    char* d = (char*)base + row_bytes * (size_t)env;
    if (to_device) CK(cudaMemcpy(d, host, bytes, cudaMemcpyHostToDevice));
    else CK(cudaMemcpy(host, d, bytes, cudaMemcpyDeviceToHost));
    return GO1_OK;
}
// diag capture is field-major [field][N]
static int copy_fields(go1_handle h, void* base, size_t elem, int nfields, int env, void* host) {
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    const size_t N = (size_t)h->n_envs;
    CK(cudaMemcpy2D(host, elem, (char*)base + elem * env, elem * N, elem, nfields, cudaMemcpyDeviceToHost));
    return GO1_OK;
}

int go1_render(go1_handle h, const Go1Camera* cam, const int32_t* env_ids_host, int n_img, int width, int height, uint8_t* rgb,
               int rgb_is_host) {
    if (!h || !env_ids_host || !rgb || n_img <= 0 || width <= 0 || height <= 0) { if (h) h->err = "go1_render: bad argument"; return GO1_EINVAL; }
    for (int i = 0; i < n_img; i++) if (env_ids_host[i] < 0 || env_ids_host[i] >= h->n_envs) { h->err = "go1_render: environment index out of range"; return GO1_EINVAL; }
#ifdef GO1_EMU
    h->err = "go1_render: not available in the CPU emulation build";
    return GO1_EUNSUPPORTED;
#else
    CK(cudaSetDevice(h->device));
    if (!h->dmodel_f32) {
        if (sizeof(real) == 4) h->dmodel_f32 = reinterpret_cast<DevModel<float>*>(h->dmodel);
        else {
            DevModel<float>* hm = new DevModel<float>();
            cast_dev_model(h->hmodel, *hm);
            cudaError_t e = cudaMalloc(&h->dmodel_f32, sizeof(DevModel<float>));
            if (e == cudaSuccess) e = cudaMemcpy(h->dmodel_f32, hm, sizeof(DevModel<float>), cudaMemcpyHostToDevice);
            delete hm;
            if (e != cudaSuccess) { h->err = std::string("go1_render: ") + cudaGetErrorString(e); return GO1_ECUDA; }
        }
    }
    const size_t bytes = (size_t)n_img * width * height * 3;
    if (n_img > h->envids_cap) { if (h->d_envids) cudaFree(h->d_envids); CK(cudaMalloc((void**)&h->d_envids, sizeof(int) * n_img)); h->envids_cap = n_img; }
    if (rgb_is_host && bytes > h->rgb_cap) { if (h->d_rgb) cudaFree(h->d_rgb); CK(cudaMalloc((void**)&h->d_rgb, bytes)); h->rgb_cap = bytes; }
    CK(cudaMemcpyAsync(h->d_envids, env_ids_host, sizeof(int) * n_img, cudaMemcpyHostToDevice, h->stream));
    RenderParams P;
    P.model = h->dmodel_f32; P.sreal = h->sreal; P.real_is_double = sizeof(real) == 8; P.s_stride = S_STRIDE; P.n_envs = h->n_envs;
    P.env_ids = h->d_envids; P.n_img = n_img; P.width = width; P.height = height;
    Go1Camera c;
    if (cam) c = *cam;
    else { const Go1Camera def = {{0.f, -1.f, 0.8f}, {1.f, 0.f, 0.f}, {0.f, 1.f, 1.f}, 45.f}; c = def; }      // go1_torque.xml:84
    P.fovy_deg = c.fovy_deg;
    // mju_makeFrame-like orthonormalisation of xyaxes; z = x cross y (the camera looks along -z)
    double x[3] = {c.xaxis[0], c.xaxis[1], c.xaxis[2]}, y[3] = {c.yaxis[0], c.yaxis[1], c.yaxis[2]};
    double nx = std::sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    if (nx < 1e-12) { h->err = "go1_render: degenerate camera axes"; return GO1_EINVAL; }
    for (int k = 0; k < 3; k++) x[k] /= nx;
    double d = x[0] * y[0] + x[1] * y[1] + x[2] * y[2];
    for (int k = 0; k < 3; k++) y[k] -= d * x[k];
    double ny = std::sqrt(y[0] * y[0] + y[1] * y[1] + y[2] * y[2]);
    if (ny < 1e-12) { h->err = "go1_render: degenerate camera axes"; return GO1_EINVAL; }
    for (int k = 0; k < 3; k++) y[k] /= ny;
    const double z[3] = {x[1] * y[2] - x[2] * y[1], x[2] * y[0] - x[0] * y[2], x[0] * y[1] - x[1] * y[0]};
    // trackcom: offset to the subtree COM of the trunk at qpos0 (mjModel.cam_poscom0)
    const DevModel<double>& M = h->hmodel;
    double com0[3] = {0, 0, 0};
    for (int k = 0; k < 3; k++) com0[k] = M.t_mass * M.t_ipos[k];
    for (int leg = 0; leg < NLEG; leg++) {
        double p[3] = {0, 0, 0};
        for (int i = 0; i < 3; i++) {
            for (int k = 0; k < 3; k++) { p[k] += M.body_pos[leg][i][k]; com0[k] += M.mass[leg][i] * (p[k] + M.ipos[leg][i][k]); }
        }
    }
    for (int k = 0; k < 3; k++) { com0[k] /= M.total_mass; P.cam_off[k] = (float)(c.pos[k] - com0[k]); P.cam_x[k] = (float)x[k]; P.cam_y[k] = (float)y[k]; P.cam_z[k] = (float)z[k]; }
    P.rgb = rgb_is_host ? h->d_rgb : rgb;
    dim3 block(16, 16), grid((width + 15) / 16, (height + 15) / 16, n_img);
    k_render<<<grid, block, 0, h->stream>>>(P);
    CK(cudaGetLastError());
    if (rgb_is_host) CK(cudaMemcpyAsync(rgb, h->d_rgb, bytes, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return GO1_OK;
#endif
}

int go1_biped_info_host(go1_handle h, float* out) {
    if (!h || !out) return GO1_EINVAL;
    if (!h->bip_info) { h->err = "go1_biped_info_host: the simulator was not created with cfg.biped"; return GO1_EUNSUPPORTED; }
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaMemcpy(out, h->bip_info, sizeof(float) * GO1_BIPED_INFO_DIM * (size_t)h->n_envs, cudaMemcpyDeviceToHost));
    return GO1_OK;
}

int go1_biped_info_dev(go1_handle h, const float** ptr) {
    if (!h || !ptr) return GO1_EINVAL;
    if (!h->bip_info) { h->err = "go1_biped_info_dev: the simulator was not created with cfg.biped"; return GO1_EUNSUPPORTED; }
    *ptr = h->bip_info;
    return GO1_OK;
}

int go1_set_state(go1_handle h, int env, const Go1State* s) {
    if (!h || !s || env < 0 || env >= h->n_envs) return GO1_EINVAL;
    real r[S_NREAL]; int iv[I_NINT];
    for (int i = 0; i < 19; i++) r[S_QPOS + i] = (real)s->qpos[i];
    for (int i = 0; i < 18; i++) { r[S_QVEL + i] = (real)s->qvel[i]; r[S_WARM + i] = (real)s->qacc_warmstart[i]; }
    for (int i = 0; i < 12; i++) r[S_LASTACT + i] = (real)s->last_action[i];
    for (int i = 0; i < 4; i++) r[S_AIR + i] = (real)s->feet_air_time[i];
    for (int i = 0; i < 3; i++) { r[S_DEV + i] = (real)s->last_health_deviation[i]; r[S_VDES + i] = (real)s->desired_velocity[i]; }
    r[S_EPRET] = (real)s->episode_return;
    iv[I_STEP] = s->step; iv[I_UNHEALTHY] = s->unhealthy_steps;
    iv[I_LASTCON] = (s->last_contacts[0] ? 1 : 0) | (s->last_contacts[1] ? 2 : 0) | (s->last_contacts[2] ? 4 : 0) | (s->last_contacts[3] ? 8 : 0);
    iv[I_RESETCNT] = s->reset_count; iv[I_BAD] = s->bad_state; iv[I_FFT] = s->front_feet_touched ? 1 : 0;
    int rc = copy_row(h, true, h->sreal, sizeof(real) * S_STRIDE, sizeof(real) * S_NREAL, env, r);
    if (rc) return rc;
    return copy_row(h, true, h->sint, sizeof(int) * I_STRIDE, sizeof(int) * I_NINT, env, iv);
}

int go1_get_state(go1_handle h, int env, Go1State* s) {
    if (!h || !s || env < 0 || env >= h->n_envs) return GO1_EINVAL;
    real r[S_NREAL]; int iv[I_NINT];
    int rc = copy_row(h, false, h->sreal, sizeof(real) * S_STRIDE, sizeof(real) * S_NREAL, env, r);
    if (rc) return rc;
    rc = copy_row(h, false, h->sint, sizeof(int) * I_STRIDE, sizeof(int) * I_NINT, env, iv);
    if (rc) return rc;
    for (int i = 0; i < 19; i++) s->qpos[i] = r[S_QPOS + i];
    for (int i = 0; i < 18; i++) { s->qvel[i] = r[S_QVEL + i]; s->qacc_warmstart[i] = r[S_WARM + i]; }
    for (int i = 0; i < 12; i++) s->last_action[i] = r[S_LASTACT + i];
    for (int i = 0; i < 4; i++) { s->feet_air_time[i] = r[S_AIR + i]; s->last_contacts[i] = (iv[I_LASTCON] >> i) & 1; }
    for (int i = 0; i < 3; i++) { s->last_health_deviation[i] = r[S_DEV + i]; s->desired_velocity[i] = r[S_VDES + i]; }
    s->episode_return = r[S_EPRET];
    s->step = iv[I_STEP]; s->unhealthy_steps = iv[I_UNHEALTHY]; s->reset_count = iv[I_RESETCNT]; s->bad_state = iv[I_BAD];
    s->front_feet_touched = iv[I_FFT]; s->pad_ = 0;
    return GO1_OK;
}

int go1_enable_diag(go1_handle h, int enable) {
    if (!h) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    if (enable && !h->diag) {
        const size_t N = (size_t)h->n_envs;
        CK(cudaMalloc((void**)&h->diag, sizeof(real) * D_NREAL * N)); CK(cudaMemset(h->diag, 0, sizeof(real) * D_NREAL * N));
        CK(cudaMalloc((void**)&h->diag_i, sizeof(int) * DI_NINT * N)); CK(cudaMemset(h->diag_i, 0, sizeof(int) * DI_NINT * N));
        const size_t cb = sizeof(real) * N * DC_ENV;
        CK(cudaMalloc((void**)&h->diag_con, cb)); CK(cudaMemset(h->diag_con, 0, cb));
    }
    h->diag_enabled = enable != 0;
    return GO1_OK;
}

int go1_get_diag(go1_handle h, int env, Go1Diag* d) {
    if (!h || !d || env < 0 || env >= h->n_envs) return GO1_EINVAL;
    if (!h->diag) { h->err = "diag capture not enabled (go1_enable_diag)"; return GO1_EINVAL; }
    real r[D_NREAL]; int iv[DI_NINT];
    int rc = copy_fields(h, h->diag, sizeof(real), D_NREAL, env, r);
    if (rc) return rc;
    rc = copy_fields(h, h->diag_i, sizeof(int), DI_NINT, env, iv);
    if (rc) return rc;
    for (int i = 0; i < 18; i++) { d->qacc[i] = r[D_QACC + i]; d->qfrc_actuator[i] = r[D_QFA + i]; d->qfrc_constraint[i] = r[D_QFC + i]; }
    for (int b = 0; b < 14; b++) for (int c = 0; c < 6; c++) d->cfrc_ext[b][c] = r[D_CFRC + 6 * b + c];
    d->trunk_xpos_z = r[D_TRUNKZ];
    for (int c = 0; c < 3; c++) d->subtree_com[c] = r[D_COM + c];
    d->ncon = iv[DI_NCON]; d->solver_iter = iv[DI_ITER]; d->solver_iter_total = iv[DI_ITERTOT]; d->contact_overflow = iv[DI_OVERFLOW];
    return GO1_OK;
}

int go1_get_contacts(go1_handle h, int env, Go1Contact* buf, int cap, int* n) {
    if (!h || !buf || !n || env < 0 || env >= h->n_envs) return GO1_EINVAL;
    if (!h->diag_con) { h->err = "diag capture not enabled (go1_enable_diag)"; return GO1_EINVAL; }
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    const size_t per = (size_t)DC_ENV;
    std::vector<real> tmp(per);
    CK(cudaMemcpy(tmp.data(), h->diag_con + (size_t)env * per, sizeof(real) * per, cudaMemcpyDeviceToHost));
    int cnt = 0;
    const int nk = (int)tmp[0];
    for (int c = 0; c < nk && c < CW && cnt < cap; c++, cnt++) {
        const real* e = tmp.data() + 1 + c * DC_STRIDE;
        buf[cnt].geom = (int)e[0] & 63; buf[cnt].geom1 = (int)e[0] >> 6; buf[cnt].pad_ = 0; buf[cnt].dim = (int)e[1]; buf[cnt].dist = e[2];
        for (int a = 0; a < 3; a++) buf[cnt].pos[a] = e[3 + a];
        for (int a = 0; a < 6; a++) buf[cnt].force[a] = e[6 + a];
    }
    *n = cnt;
    return GO1_OK;
}

int go1_set_attr_f32(go1_handle h, const char* name, float v) {
    if (!h || !name) return GO1_EINVAL;
    std::string s(name);
    if (s == "rand_power" || s == "_rand_power") h->cfg.rand_power = v;
    else if (s == "fallen_start_prob") h->cfg.fallen_prob = v;
    else if (s.rfind("desired_vel_min_", 0) == 0 && s.size() == 17 && s[16] >= 'x' && s[16] <= 'z') h->cfg.vdes_min[s[16] - 'x'] = v;
    else if (s.rfind("desired_vel_max_", 0) == 0 && s.size() == 17 && s[16] >= 'x' && s[16] <= 'z') h->cfg.vdes_max[s[16] - 'x'] = v;
    else if (s == "solver_tolerance") h->cfg.tolerance = (real)v;
    else { h->err = "unknown attribute " + s; return GO1_EINVAL; }
    return GO1_OK;
}

int go1_get_attr_f32(go1_handle h, const char* name, float* v) {
    if (!h || !name || !v) return GO1_EINVAL;
    std::string s(name);
    if (s == "rand_power" || s == "_rand_power") *v = (float)h->cfg.rand_power;
    else if (s == "fallen_start_prob") *v = (float)h->cfg.fallen_prob;
    else if (s.rfind("desired_vel_min_", 0) == 0 && s.size() == 17 && s[16] >= 'x' && s[16] <= 'z') *v = (float)h->cfg.vdes_min[s[16] - 'x'];
    else if (s.rfind("desired_vel_max_", 0) == 0 && s.size() == 17 && s[16] >= 'x' && s[16] <= 'z') *v = (float)h->cfg.vdes_max[s[16] - 'x'];
    else if (s == "solver_tolerance") *v = (float)h->cfg.tolerance;
    else { h->err = "unknown attribute " + s; return GO1_EINVAL; }
    return GO1_OK;
}

}  // extern "C"

#ifdef GO1_EMU_COUNT
long g_emu_rounds[4] = {0, 0, 0, 0};      // test-only instrumentation of the emulator build: Newton rounds by Hessian path
extern "C" void go1_emu_round_counts(long* out) { for (int i = 0; i < 4; i++) out[i] = g_emu_rounds[i]; }
#endif

#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
// diagnostic build only: reads and clears the per-class phase clocks (tools/phase_times.py)
extern "C" int go1_debug_phase_times(unsigned long long* out80) {
    if (cudaMemcpyFromSymbol(out80, go1::g_phase, sizeof(unsigned long long) * 80) != cudaSuccess) return GO1_ECUDA;
    unsigned long long z[80] = {};
    return cudaMemcpyToSymbol(go1::g_phase, z, sizeof(z)) == cudaSuccess ? GO1_OK : GO1_ECUDA;
}
#endif

