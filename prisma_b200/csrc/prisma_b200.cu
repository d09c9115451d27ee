/* prisma_b200 — sm_100a kernels and the C ABI of include/prisma_b200.h.
 *
 * Execution model: one group of G lanes (default G = 32, a full warp) owns one environment.  The
 * environment's state is one contiguous blob in HBM (layout: model.h) plus a packet-record pool.  Per
 * simulated event every lane scans its share of the pending-timer slots, the group agrees on the
 * minimum (timestamp, scheduler uid) with warp reductions, and the group leader executes the event
 * (sim_core.cuh).  In run mode the blob is staged in shared memory for the whole launch, so HBM sees one
 * read and one write of the blob per launch plus the packet records that move.
 *
 * There is no host execution path in this file: every entry point that advances an environment launches
 * a kernel.
 */
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "model.h"

using namespace prb;

#define PRB_VERSION "prisma_b200 0.1 (sm_100a)"

static thread_local std::string g_err;

static int
fail (int code, const std::string &msg)
{
  g_err = msg;
  return code;
}

#define CUDA_TRY(expr)                                                                                   \
  do                                                                                                     \
    {                                                                                                    \
      cudaError_t e_ = (expr);                                                                           \
      if (e_ != cudaSuccess)                                                                             \
        return fail (PRB_ECUDA, std::string (#expr) + ": " + cudaGetErrorString (e_));                   \
    }                                                                                                    \
  while (0)

/* ------------------------------------------------------------------------------------------------
 * Device side
 * ---------------------------------------------------------------------------------------------- */

/* ------------------------------------------------------------------------------------------------
 * Batched Q-network of the dqn_buffer agents (prisma/source/models.py:257-301, learner.py:142-159):
 * one network per overlay node,
 *     h_d = elu (W_dst[dst] + b_dst)                      one-hot destination -> Dense(H1): a row lookup
 *     h_b = elu (LayerNorm (buffers) W_buf + b_buf)       LayerNormalization without affine, epsilon 1e-3
 *     h_1 = elu ([h_d, h_b] W_1 + b_1)   h_2 = elu (h_1 W_2 + b_2)   q = elu (h_2 W_3 + b_3)
 *     action = argmin q, or a uniform random neighbour with probability eps (DQN_AGENT.step).
 * One warp per environment evaluates the pending Gym step of that environment (prb_step_rec) and writes the
 * action prb_step consumes next.  17.7 kFLOP per decision, 70 MFLOP for 4096 environments: the step is launch
 * bound, so this is a CUDA-core kernel (fp32, fixed summation order) and not a tensor-core GEMM.
 * Weights of node m (floats, Keras kernel layout [in][out]):
 *     W_dst[M][H1] b_dst[H1] W_buf[K][H1] b_buf[H1] W_1[2*H1][H2] b_1[H2] W_2[H2][H2] b_2[H2] W_3[H2][K] b_3[K]
 * with H1 = 32, H2 = 64. ---------------------------------------------------------------------------- */
constexpr int QH1 = 32, QH2 = 64;

__host__ __device__ inline int64_t
qnet_floats_per_node (int M, int K)
{
  return (int64_t) M * QH1 + QH1 + (int64_t) K * QH1 + QH1 + 2 * QH1 * QH2 + QH2 + QH2 * QH2 + QH2 + (int64_t) QH2 * K + K;
}

#ifndef PRB_ELU_ATTR
#define PRB_ELU_ATTR __forceinline__
#endif
__device__ PRB_ELU_ATTR float
elu_ (float x)
{
  return x > 0.0f ? x : expm1f (x);
}

/* all 32 lanes of a warp: action for the pending data step *r of environment-id eid (every lane returns it).
 * One copy in the instruction stream and modest unrolling (4): fully unrolled and inlined at both call
 * sites of k_advance it was 8 268 SASS instructions, and the 27 warps of an SM, each at its own point of a 50 KB hot
 * loop, spent 18 of 25 cycles per issue waiting for instructions (profiles/r1_k_advance_gym_*).  The order of the
 * additions of every accumulator is the loop order either way. */
#ifndef PRB_QNET_ATTR
#define PRB_QNET_ATTR __forceinline__
#endif
#ifndef PRB_QNET_UNROLL
#define PRB_QNET_UNROLL 8
#endif
constexpr int QNET_UNROLL = PRB_QNET_UNROLL;
template <bool ONE_HIDDEN = false>
__device__ PRB_QNET_ATTR int
qnet_forward (const Static &st, const float *w, const prb_step_rec *r, int lane, float eps, uint64_t seed, uint64_t step, uint64_t eid,
              float *q_row, float *tmin = nullptr)
{
  const int m = r->overlay, K = st.K, M = st.M;
  const int k = st.agent_k[m];
  const float *W = w + (size_t) m * qnet_floats_per_node (M, K);
  const float *w_dst = W, *b_dst = w_dst + (size_t) M * QH1, *w_buf = b_dst + QH1, *b_buf = w_buf + (size_t) K * QH1;
  const float *w1 = b_buf + QH1, *b1 = w1 + 2 * QH1 * QH2, *w2 = b1 + QH2, *b2 = w2 + QH2 * QH2, *w3 = b2 + QH2, *b3 = w3 + (size_t) QH2 * K;
  /* destination branch: row obs[0] of W_dst */
  int dst = r->obs[0];
  dst = dst < 0 ? 0 : (dst >= M ? M - 1 : dst);
  float hd = elu_ (w_dst[(size_t) dst * QH1 + lane] + b_dst[lane]);
  /* buffer branch: layer normalisation over the k observed queues, then Dense(H1) */
  float x = lane < k ? (float) (uint32_t) r->obs[1 + lane] : 0.0f;
  float sum = x;
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync (0xFFFFFFFFu, sum, o);
  float mean = sum / (float) k;
  float dv = lane < k ? x - mean : 0.0f;
  float var = dv * dv;
  for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync (0xFFFFFFFFu, var, o);
  float xn = dv * rsqrtf (var / (float) k + 1e-3f);
  float acc = b_buf[lane];
  for (int i = 0; i < k; i++) acc = fmaf (__shfl_sync (0xFFFFFFFFu, xn, i), w_buf[(size_t) i * QH1 + lane], acc);
  float hb = elu_ (acc);
  /* Dense(64) on [h_d, h_b]: lane j owns outputs j and j + 32 */
  float a0 = b1[lane], a1 = b1[lane + 32];
#pragma unroll QNET_UNROLL
  for (int i = 0; i < QH1; i++)
    {
      float c = __shfl_sync (0xFFFFFFFFu, hd, i);
      a0 = fmaf (c, w1[(size_t) i * QH2 + lane], a0);
      a1 = fmaf (c, w1[(size_t) i * QH2 + lane + 32], a1);
    }
#pragma unroll QNET_UNROLL
  for (int i = 0; i < QH1; i++)
    {
      float c = __shfl_sync (0xFFFFFFFFu, hb, i);
      a0 = fmaf (c, w1[(size_t) (QH1 + i) * QH2 + lane], a0);
      a1 = fmaf (c, w1[(size_t) (QH1 + i) * QH2 + lane + 32], a1);
    }
  float h0 = elu_ (a0), h1 = elu_ (a1);
  if (!ONE_HIDDEN)             /* the one-hidden-layer models (lighter*) go from here straight to the output layer */
    {
      a0 = b2[lane]; a1 = b2[lane + 32];
#pragma unroll QNET_UNROLL
      for (int i = 0; i < 32; i++)
        {
          float c = __shfl_sync (0xFFFFFFFFu, h0, i);
          a0 = fmaf (c, w2[(size_t) i * QH2 + lane], a0);
          a1 = fmaf (c, w2[(size_t) i * QH2 + lane + 32], a1);
        }
#pragma unroll QNET_UNROLL
      for (int i = 0; i < 32; i++)
        {
          float c = __shfl_sync (0xFFFFFFFFu, h1, i);
          a0 = fmaf (c, w2[(size_t) (32 + i) * QH2 + lane], a0);
          a1 = fmaf (c, w2[(size_t) (32 + i) * QH2 + lane + 32], a1);
        }
      h0 = elu_ (a0); h1 = elu_ (a1);
    }
  /* Dense(k): lane j < k owns q_j */
  float q = lane < k ? b3[lane] : 0.0f;
#pragma unroll QNET_UNROLL
  for (int i = 0; i < 32; i++)
    {
      float c = __shfl_sync (0xFFFFFFFFu, h0, i);
      if (lane < k) q = fmaf (c, w3[(size_t) i * K + lane], q);
    }
#pragma unroll QNET_UNROLL
  for (int i = 0; i < 32; i++)
    {
      float c = __shfl_sync (0xFFFFFFFFu, h1, i);
      if (lane < k) q = fmaf (c, w3[(size_t) (32 + i) * K + lane], q);
    }
  q = elu_ (q);
  if (q_row && lane < K) q_row[lane] = lane < k ? q : 0.0f;
  if (tmin)
    {
      /* `target` signalling: what this node reports back to the previous decision maker is the smallest Q-value
       * over its actions except the one that leads back there (trainer.py:58-60 filtered_indices,
       * learner.py:232-255); +inf when no action is left */
      float tq = (lane < k && (int) st.agent_nbr[(size_t) m * K + lane] != r->prev_node) ? q : INFINITY;
      for (int o = 16; o > 0; o >>= 1) tq = fminf (tq, __shfl_xor_sync (0xFFFFFFFFu, tq, o));
      *tmin = tq;
    }
  /* argmin, first index on equal values (tf.argmin) */
  float bq = lane < k ? q : INFINITY;
  int bi = lane;
  for (int o = 16; o > 0; o >>= 1)
    {
      float oq = __shfl_xor_sync (0xFFFFFFFFu, bq, o);
      int oi = __shfl_xor_sync (0xFFFFFFFFu, bi, o);
      if (oq < bq || (oq == bq && oi < bi))
        {
          bq = oq; bi = oi;
        }
    }
  int action = bi;
  if (eps > 0.0f)
    {
      uint32_t rr[4];
      philox4x32 ((uint32_t) eid, (uint32_t) (eid >> 32), (uint32_t) step, (uint32_t) (step >> 32), (uint32_t) seed,
                  (uint32_t) (seed >> 32) ^ 0x51DEu, rr);
      float u = (float) (rr[0] >> 8) * (1.0f / 16777216.0f);
      if (u < eps) action = (int) (rr[1] % (uint32_t) k);
    }
  return action;
}

__global__ void __launch_bounds__ (128)
k_qnet_act (const __grid_constant__ Static st, const float *w, const prb_step_rec *rec, int32_t *actions, float *q_out,
            float eps, uint64_t seed, uint64_t step, float *tmin_out)
{
  const int lane = threadIdx.x & 31;
  const int e = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  if (e >= st.E) return;
  const prb_step_rec *r = &rec[e];
  if (!r->valid || !r->need_action)
    {
      if (lane == 0) actions[e] = 0;
      return;
    }
  float tmin = INFINITY;
  float *qr = q_out ? q_out + (size_t) e * st.K : nullptr;
  int action = st.q_hidden > 1 ? qnet_forward<false> (st, w, r, lane, eps, seed, step, st.env_id_base + (uint64_t) e, qr, tmin_out ? &tmin : nullptr)
                               : qnet_forward<true> (st, w, r, lane, eps, seed, step, st.env_id_base + (uint64_t) e, qr, tmin_out ? &tmin : nullptr);
  if (lane == 0)
    {
      actions[e] = action;
      if (tmin_out) tmin_out[e] = tmin;
    }
}

/* ------------------------------------------------------------------------------------------------
 * Replay buffers in HBM (prisma/source/replay_buffer.py:12-86), fed from the step records the way Forwarder
 * does it (prisma/source/forwarder.py:135-161 temp_obs, :214-244 lost packets, :352-379 transit packets with
 * 'ideal' signalling).  One ring per overlay node shared by all environments; a row is
 *     obs[1+K] action reward next_obs[1+K] done            (float32, the dtype the reference stores: float)
 * and a per-environment table keyed by packet uid holds the decision a packet is travelling with
 * (Agent.temp_obs).  One thread per environment: the record of an environment is one packet. */
struct ReplayArgs
{
  float *rows;                /* [M][cap][R] */
  unsigned long long *cursor; /* [M] rows ever written */
  unsigned long long *missed; /* transitions whose decision was not found (table collision / lost list overflow) */
  int32_t *pending;           /* [E][pcap][PW]: uid lo, uid hi, node, action, obs[1+K] */
  int64_t cap;
  int32_t pcap, R, PW;
  /* `target` signalling (trainer.py:52-55): the reward column holds the TD target the NEXT node computes with its
   * own network when the packet reaches it: hop time + gamma * min over its actions that do not lead back */
  int32_t target_mode;
  float gamma;
  float *tmin;                /* [E]: that minimum for the pending step of every environment (unfused path) */
  /* prioritized variant (replay_buffer.py:393-534): a row remembers the gradient step of the network that produced
   * its target, a node the latest gradient step it has seen per action; weight = (row / latest) ^ alpha */
  float *row_gstep;           /* [M][cap] or null */
  int32_t *latest;            /* [M][K] */
  const int32_t *node_gstep;  /* [M] gradient steps taken by every node's learner (host-set) */
  double *psum;               /* scratch: sum of the priorities of one node */
  float alpha;
};

__device__ inline void
replay_push (const Static &st, const ReplayArgs &ra, int node, const int32_t *ent, float reward, const int32_t *next_obs,
             int next_n, int done, int from_node = -1)
{
  unsigned long long slot = atomicAdd (&ra.cursor[node], 1ull) % (unsigned long long) ra.cap;
  if (ra.row_gstep)
    {
      /* forwarder.py:494-505: a transit transition carries the gradient step of the node that computed its target
       * and raises the latest step seen for that action; :227-233: a lost packet is added with the latest one */
      const int act = ent[3] < 0 ? 0 : (ent[3] >= st.K ? st.K - 1 : ent[3]);    /* rows only ever hold valid actions */
      int32_t *lat = &ra.latest[(size_t) node * st.K + act];
      int32_t g = from_node >= 0 ? ra.node_gstep[from_node] : *lat;
      if (from_node >= 0) atomicMax (lat, g);
      ra.row_gstep[(size_t) node * ra.cap + slot] = (float) g;
    }
  float *row = ra.rows + ((size_t) node * ra.cap + slot) * ra.R;
  const int ow = 1 + st.K;
  for (int i = 0; i < ow; i++) row[i] = (float) ent[4 + i];
  row[ow] = (float) ent[3];
  row[ow + 1] = reward;
  for (int i = 0; i < ow; i++) row[ow + 2 + i] = i < next_n ? (float) next_obs[i] : 0.0f;
  row[2 * ow + 2] = (float) done;
}

/* the bookkeeping of one step record (one thread) */
__device__ inline void
replay_record_one (const Static &st, const ReplayArgs &ra, int e, const prb_step_rec *r, int32_t action, float tmin = INFINITY)
{
  if (!r->valid || r->ptype != PRB_DATA) return;
  int32_t *tab = ra.pending + (size_t) e * ra.pcap * ra.PW;
  const int ow = 1 + st.K;
  /* lost packets reported with this step: (s, a, loss_penalty, [dst, 0...], done) into THIS node's buffer */
  int nl = r->n_lost;
  if (nl > PRB_MAX_LOST)
    {
      atomicAdd (ra.missed, (unsigned long long) (nl - PRB_MAX_LOST));
      nl = PRB_MAX_LOST;
    }
  for (int j = 0; j < nl; j++)
    {
      int64_t uid = r->lost_uid[j];
      int32_t *ent = tab + (size_t) (uid & (ra.pcap - 1)) * ra.PW;
      if (ent[0] == (int32_t) uid && ent[1] == (int32_t) (uid >> 32) && ent[2] >= 0)
        {
          int32_t nxt[1] = { ent[4] };
          replay_push (st, ra, r->overlay, ent, st.loss_penalty, nxt, 1, 1);
          ent[2] = -1;
        }
      else
        atomicAdd (ra.missed, 1ull);
    }
  int32_t *ent = tab + (size_t) (r->uid & (ra.pcap - 1)) * ra.PW;
  /* a packet that travelled from another agent: the transition of that agent's decision */
  if (!r->is_new)
    {
      if (ent[0] == (int32_t) r->uid && ent[1] == (int32_t) (r->uid >> 32) && ent[2] >= 0)
        {
          float rew = (float) r->hop_time;
          if (ra.target_mode && r->need_action && tmin < 1e37f) rew = rew + ra.gamma * tmin;
          replay_push (st, ra, ent[2], ent, rew, r->obs, r->nobs, r->done, r->overlay);
          ent[2] = -1;
        }
      else
        atomicAdd (ra.missed, 1ull);
    }
  /* the decision taken now travels with the packet — unless the action names no neighbour: the packet vanishes
   * (data-packet-manager.cc:294-297), the reference never completes such a transition, and rows index per-action
   * tables with it */
  if (r->need_action && (action < 0 || action >= (int32_t) st.agent_k[r->overlay]))
    {
      atomicAdd (ra.missed, 1ull);
      if (ent[0] == (int32_t) r->uid && ent[1] == (int32_t) (r->uid >> 32)) ent[2] = -1;
    }
  else if (r->need_action)
    {
      ent[0] = (int32_t) r->uid; ent[1] = (int32_t) (r->uid >> 32); ent[2] = r->overlay; ent[3] = action;
      for (int i = 0; i < ow; i++) ent[4 + i] = r->obs[i];
    }
}

__global__ void
k_replay_record (const __grid_constant__ Static st, const ReplayArgs ra, const prb_step_rec *rec, const int32_t *actions)
{
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= st.E) return;
  replay_record_one (st, ra, e, &rec[e], actions[e], ra.target_mode && ra.tmin ? ra.tmin[e] : INFINITY);
}

/* ------------------------------------------------------------------------------------------------
 * Batched Q-network forward on the tensor cores (DQN_buffer_model, prisma/source/models.py:258-306, and
 * DQN_AGENT.step, prisma/source/learner.py:142-159, for the pending decisions of ALL environments at once).
 * k_qnet_bucket sorts the pending decisions by overlay node (one atomic per decision); k_qnet_mma gives every warp a
 * tile of 16 decisions of ONE node and evaluates Dense64 -> Dense64 -> Dense(k) as mma.sync.m16n8k8 TF32 tiles with fp32
 * accumulators.  Every product is taken three times on split operands (x = hi + lo, hi * hi + lo * hi + hi * lo: "3xTF32"),
 * which brings the result back to fp32 accuracy (the reference's TensorFlow runs these layers in fp32 or TF32 depending on
 * the GPU; the fused fp32 path of k_advance stays the numerical yardstick of the tests).  Why mma.sync and not
 * tcgen05: one call is 70 MFLOP (4096 decisions x 17.7 kFLOP) in tiles of 16 x 64 x 64 — a tcgen05 tile is at least
 * 64 rows of one node, TMEM allocation and descriptor set-up would outlast the arithmetic, and the kernel is bound by the
 * launch and by the latency of the weight loads either way.  The destination branch (a row lookup) and the buffer branch
 * (LayerNorm + Dense of k <= 16 inputs) are fp32 FMAs.  After the last layer one lane per decision takes the argmin /
 * eps-greedy draw (same Philox stream as qnet_forward), writes the action and feeds the replay rings.
 * ---------------------------------------------------------------------------------------------- */
struct QBatchArgs
{
  int32_t *count;          /* [2][M + 1]: decisions per node of this call (parity: the other half is cleared for the next call) */
  int32_t *perm;           /* [M][E]: environment indices per node */
  int parity;
};

__global__ void
k_qnet_bucket (const __grid_constant__ Static st, const ReplayArgs ra, const prb_step_rec *rec, int32_t *actions, QBatchArgs qb)
{
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= st.E) return;
  const prb_step_rec *r = &rec[e];
  if (!r->valid || !r->need_action)
    {
      actions[e] = 0;
      /* a step without a decision (the packet arrived) still closes the previous node's transition */
      if (ra.rows) replay_record_one (st, ra, e, r, 0, INFINITY);
      return;
    }
  const int m = r->overlay;
  const int pos = atomicAdd (&qb.count[qb.parity * (st.M + 1) + m], 1);
  qb.perm[(size_t) m * st.E + pos] = e;
}

__device__ __forceinline__ uint32_t
tf32_hi (float x)
{
  uint32_t r;
  asm ("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

__device__ __forceinline__ void
mma_tf32 (float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
  asm volatile ("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

/* c += a * b with a (16 x 8 of the activations) and b (8 x 8 of the weights) split into TF32 hi + lo parts */
__device__ __forceinline__ void
mma_3xtf32 (float (&c)[4], const float (&a)[4], float b0, float b1)
{
  uint32_t ah[4], al[4];
#pragma unroll
  for (int i = 0; i < 4; i++)
    {
      ah[i] = tf32_hi (a[i]);
      al[i] = tf32_hi (a[i] - __uint_as_float (ah[i]));
    }
  const uint32_t bh0 = tf32_hi (b0), bh1 = tf32_hi (b1);
  const uint32_t bl0 = tf32_hi (b0 - __uint_as_float (bh0)), bl1 = tf32_hi (b1 - __uint_as_float (bh1));
  mma_tf32 (c, al, bh0, bh1);
  mma_tf32 (c, ah, bl0, bl1);
  mma_tf32 (c, ah, bh0, bh1);
}

constexpr int QM_WARPS = 2;          /* warps (tiles of 16 decisions) per block */
constexpr int QM_LD = QH2 + 4;       /* row pitch of the activation tile in shared memory (floats): conflict-free fragment reads */

/* one Dense(64 -> 64) + elu on a 16-row tile: activations in act (shared, [16][QM_LD]), result written back in place */
__device__ __forceinline__ void
qm_dense64 (float *act, const float *__restrict__ w, const float *__restrict__ b, int g, int t)
{
  float c[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; nt++)
    {
      const float b0 = b[nt * 8 + 2 * t], b1 = b[nt * 8 + 2 * t + 1];
      c[nt][0] = b0; c[nt][1] = b1; c[nt][2] = b0; c[nt][3] = b1;
    }
#pragma unroll
  for (int kk = 0; kk < 8; kk++)
    {
      float a[4];
      a[0] = act[g * QM_LD + kk * 8 + t];
      a[1] = act[(g + 8) * QM_LD + kk * 8 + t];
      a[2] = act[g * QM_LD + kk * 8 + t + 4];
      a[3] = act[(g + 8) * QM_LD + kk * 8 + t + 4];
#pragma unroll
      for (int nt = 0; nt < 8; nt++)
        mma_3xtf32 (c[nt], a, __ldg (&w[(size_t) (kk * 8 + t) * QH2 + nt * 8 + g]), __ldg (&w[(size_t) (kk * 8 + t + 4) * QH2 + nt * 8 + g]));
    }
  __syncwarp ();
#pragma unroll
  for (int nt = 0; nt < 8; nt++)
    {
      act[g * QM_LD + nt * 8 + 2 * t] = elu_ (c[nt][0]);
      act[g * QM_LD + nt * 8 + 2 * t + 1] = elu_ (c[nt][1]);
      act[(g + 8) * QM_LD + nt * 8 + 2 * t] = elu_ (c[nt][2]);
      act[(g + 8) * QM_LD + nt * 8 + 2 * t + 1] = elu_ (c[nt][3]);
    }
  __syncwarp ();
}

__global__ void __launch_bounds__ (QM_WARPS * 32)
k_qnet_mma (const __grid_constant__ Static st, const ReplayArgs ra, const float *__restrict__ wts, const prb_step_rec *rec,
            int32_t *actions, float *q_out, float eps, uint64_t seed, uint64_t step, float *tmin_out, QBatchArgs qb)
{
  __shared__ __align__ (16) float s_act[QM_WARPS][16 * QM_LD];
  __shared__ float s_q[QM_WARPS][16 * PRB_MAX_K];
  __shared__ int s_env[QM_WARPS][16];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int M = st.M, K = st.K;
  const int32_t *count = qb.count + qb.parity * (M + 1);
  /* the other half of the counters is the next call's: clear it */
  if (blockIdx.x == 0 && threadIdx.x <= M) qb.count[(qb.parity ^ 1) * (M + 1) + threadIdx.x] = 0;
  /* which node, which 16 decisions of it */
  int tile = blockIdx.x * QM_WARPS + warp, m = 0, n_m = 0;
  for (; m < M; m++)
    {
      n_m = count[m];
      const int tiles = (n_m + 15) >> 4;
      if (tile < tiles) break;
      tile -= tiles;
    }
  if (m >= M) return;
  const int row0 = tile * 16;
  const int k = st.agent_k[m];
  const float *W = wts + (size_t) m * qnet_floats_per_node (M, K);
  const float *w_dst = W, *b_dst = w_dst + (size_t) M * QH1, *w_buf = b_dst + QH1, *b_buf = w_buf + (size_t) K * QH1;
  const float *w1 = b_buf + QH1, *b1 = w1 + 2 * QH1 * QH2, *w2 = b1 + QH2, *b2 = w2 + QH2 * QH2, *w3 = b2 + QH2, *b3 = w3 + (size_t) QH2 * K;
  float *act = s_act[warp];
  if (lane < 16) s_env[warp][lane] = row0 + lane < n_m ? qb.perm[(size_t) m * st.E + row0 + lane] : -1;
  __syncwarp ();
  /* input branches, fp32: lane j owns unit j of both; the k queue values are normalised by every lane alike */
  for (int r = 0; r < 16; r++)
    {
      const int e = s_env[warp][r];
      float hd = 0.0f, hb = 0.0f;
      if (e >= 0)
        {
          const prb_step_rec *rc = &rec[e];
          int dst = rc->obs[0];
          dst = dst < 0 ? 0 : (dst >= M ? M - 1 : dst);
          hd = elu_ (w_dst[(size_t) dst * QH1 + lane] + b_dst[lane]);
          float sum = 0.0f;
          for (int i = 0; i < k; i++) sum += (float) (uint32_t) rc->obs[1 + i];
          const float mean = sum / (float) k;
          float var = 0.0f;
          for (int i = 0; i < k; i++)
            {
              const float dv = (float) (uint32_t) rc->obs[1 + i] - mean;
              var += dv * dv;
            }
          const float rs = rsqrtf (var / (float) k + 1e-3f);
          float a = b_buf[lane];
          for (int i = 0; i < k; i++) a = fmaf (((float) (uint32_t) rc->obs[1 + i] - mean) * rs, w_buf[(size_t) i * QH1 + lane], a);
          hb = elu_ (a);
        }
      act[r * QM_LD + lane] = hd;
      act[r * QM_LD + QH1 + lane] = hb;
    }
  __syncwarp ();
  qm_dense64 (act, w1, b1, g, t);
  if (st.q_hidden > 1) qm_dense64 (act, w2, b2, g, t);
  /* Dense(k): columns >= k are padding */
  for (int nt = 0; nt * 8 < K; nt++)
    {
      float c[4];
      const int n0 = nt * 8 + 2 * t, n1 = n0 + 1;
      c[0] = c[2] = n0 < k ? b3[n0] : 0.0f;
      c[1] = c[3] = n1 < k ? b3[n1] : 0.0f;
      const int nb = nt * 8 + g;
#pragma unroll
      for (int kk = 0; kk < 8; kk++)
        {
          float a[4];
          a[0] = act[g * QM_LD + kk * 8 + t];
          a[1] = act[(g + 8) * QM_LD + kk * 8 + t];
          a[2] = act[g * QM_LD + kk * 8 + t + 4];
          a[3] = act[(g + 8) * QM_LD + kk * 8 + t + 4];
          mma_3xtf32 (c, a, nb < k ? __ldg (&w3[(size_t) (kk * 8 + t) * K + nb]) : 0.0f,
                      nb < k ? __ldg (&w3[(size_t) (kk * 8 + t + 4) * K + nb]) : 0.0f);
        }
      if (n0 < K)
        {
          s_q[warp][g * PRB_MAX_K + n0] = n0 < k ? elu_ (c[0]) : 0.0f;
          s_q[warp][(g + 8) * PRB_MAX_K + n0] = n0 < k ? elu_ (c[2]) : 0.0f;
        }
      if (n1 < K)
        {
          s_q[warp][g * PRB_MAX_K + n1] = n1 < k ? elu_ (c[1]) : 0.0f;
          s_q[warp][(g + 8) * PRB_MAX_K + n1] = n1 < k ? elu_ (c[3]) : 0.0f;
        }
    }
  __syncwarp ();
  /* one lane per decision: argmin (first index on equal values, tf.argmin), eps-greedy, replay bookkeeping */
  if (lane < 16 && s_env[warp][lane] >= 0)
    {
      const int e = s_env[warp][lane];
      const prb_step_rec *rc = &rec[e];
      const float *q = &s_q[warp][lane * PRB_MAX_K];
      float bq = INFINITY, tq = INFINITY;
      int action = 0;
      for (int i = 0; i < k; i++)
        {
          if (q[i] < bq)
            {
              bq = q[i]; action = i;
            }
          if ((int) st.agent_nbr[(size_t) m * K + i] != rc->prev_node) tq = fminf (tq, q[i]);
        }
      if (q_out)
        for (int i = 0; i < K; i++) q_out[(size_t) e * K + i] = q[i];
      if (eps > 0.0f)
        {
          const uint64_t eid = st.env_id_base + (uint64_t) e;
          uint32_t rr[4];
          philox4x32 ((uint32_t) eid, (uint32_t) (eid >> 32), (uint32_t) step, (uint32_t) (step >> 32), (uint32_t) seed,
                      (uint32_t) (seed >> 32) ^ 0x51DEu, rr);
          const float u = (float) (rr[0] >> 8) * (1.0f / 16777216.0f);
          if (u < eps) action = (int) (rr[1] % (uint32_t) k);
        }
      actions[e] = action;
      if (tmin_out) tmin_out[e] = tq;
      if (ra.rows) replay_record_one (st, ra, e, rc, action, ra.target_mode ? tq : INFINITY);
    }
}

#include "learner_core.cuh"

/* priority of row idx of a node: (gradient step of the row / latest step seen for its action) ^ alpha */
__device__ __forceinline__ float
replay_priority (const ReplayArgs &ra, int node, unsigned long long idx, int ow, int K)
{
  const float *row = ra.rows + ((size_t) node * ra.cap + idx) * ra.R;
  int a = (int) row[ow];
  a = a < 0 ? 0 : (a >= K ? K - 1 : a);
  float lat = (float) ra.latest[(size_t) node * K + a];
  float g = ra.row_gstep[(size_t) node * ra.cap + idx];
  return powf (g / (lat > 0.0f ? lat : 1.0f), ra.alpha);
}

__global__ void
k_replay_priority_sum (const ReplayArgs ra, int node, int ow, int K)
{
  unsigned long long n = ra.cursor[node];
  if (n > (unsigned long long) ra.cap) n = (unsigned long long) ra.cap;
  double acc = 0.0;
  for (unsigned long long i = (unsigned long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long) gridDim.x * blockDim.x)
    acc += (double) replay_priority (ra, node, i, ow, K);
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync (0xFFFFFFFFu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc != 0.0) atomicAdd (ra.psum, acc);
}

__global__ void
k_replay_sample (const ReplayArgs ra, int node, int batch, int ow, uint64_t seed, uint64_t call, float *obs, int32_t *act,
                 float *rew, float *next_obs, float *done, float *weights = nullptr, int K = 0)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  unsigned long long n = ra.cursor[node];
  if (n > (unsigned long long) ra.cap) n = (unsigned long long) ra.cap;
  if (n == 0) return;
  uint32_t rr[4];
  philox4x32 ((uint32_t) b, (uint32_t) node, (uint32_t) call, (uint32_t) (call >> 32), (uint32_t) seed, (uint32_t) (seed >> 32) ^ 0x5A3Du, rr);
  unsigned long long idx = (((unsigned long long) rr[0] << 32) | rr[1]) % n;     /* random.randint(0, len - 1) */
  const float *row = ra.rows + ((size_t) node * ra.cap + idx) * ra.R;
  for (int i = 0; i < ow; i++) obs[(size_t) b * ow + i] = row[i];
  act[b] = (int32_t) row[ow];
  rew[b] = row[ow + 1];
  for (int i = 0; i < ow; i++) next_obs[(size_t) b * ow + i] = row[ow + 2 + i];
  done[b] = row[2 * ow + 2];
  /* PrioritizedReplayBuffer.sample: uniform draws, weight = priority / sum of priorities * len (replay_buffer.py:497-498) */
  if (weights) weights[b] = ra.row_gstep ? (float) ((double) replay_priority (ra, node, idx, ow, K) / *ra.psum * (double) n) : 1.0f;
}


struct LaunchArgs
{
  char *blobs;            /* [E * blob_bytes] */
  Pkt *pools;             /* [E * P] */
  prb_event *ev_logs;     /* [E * ev_cap] or null */
  prb_step_rec *st_logs;  /* [E * st_cap] or null */
  const int32_t *actions; /* gym mode, may be null */
  prb_step_rec *out;      /* gym mode */
  int64_t until_ns;
  /* gym mode with the dqn_buffer agents evaluated inside the kernel (prb_step_qnet): every environment answers its
   * own steps until it has taken `steps` of them */
  const float *qw;        /* null: one step per launch, actions from `actions` */
  int32_t steps;
  float eps;
  uint64_t seed, step0;
  ReplayArgs replay;      /* replay.rows == null: no replay buffers */
  /* run mode on a subset: slot i of the grid advances environment env_list[i] (the parallel engine's hand-overs) */
  const int32_t *env_list;
  int32_t n_list;
};

/* minimum (key, uid) over the lanes of a group; every lane gets the winning slot (or -1) and its key */
template <int G>
__device__ __forceinline__ void
group_min (unsigned gmask, const uint32_t *t_seq, uint32_t key, int slot, uint32_t &mkey, int &mslot)
{
  mkey = __reduce_min_sync (gmask, key);
  bool cand = key == mkey;
  unsigned b = __ballot_sync (gmask, cand);
  if (mkey != KEY_IDLE && (b & (b - 1)))
    {
      /* several lanes hold the same timestamp: the scheduler uid decides */
      uint32_t q = cand ? t_seq[slot] : 0xFFFFFFFFu;
      uint32_t mq = __reduce_min_sync (gmask, q);
      b = __ballot_sync (gmask, cand && q == mq);
    }
  mslot = __shfl_sync (gmask, slot, __ffs (b) - 1);
}

/* the earliest timer of flow group g, found by the G lanes of the environment together (sim_core.cuh: scan_flow_groups) */
template <int G>
__device__ __forceinline__ void
fg_rescan_coop (const Static &st, const Dyn &d, int g, int gl, unsigned gmask)
{
  int64_t bt = T_IDLE; uint32_t bq = 0xFFFFFFFFu; int bs = -1;
  const int s0 = st.NK + 32 * g, s1 = s0 + 32 < st.S ? s0 + 32 : st.S;
  for (int s = s0 + gl; s < s1; s += G)
    {
      int64_t t = d.t_ts[s];
      if (t == T_IDLE || t > bt) continue;
      uint32_t q = d.t_seq[s];
      if (t < bt || q < bq)
        {
          bt = t; bq = q; bs = s;
        }
    }
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1)
    {
      const int64_t ot = __shfl_xor_sync (gmask, bt, o);
      const uint32_t oq = __shfl_xor_sync (gmask, bq, o);
      const int os = __shfl_xor_sync (gmask, bs, o);
      if (ot < bt || (ot == bt && oq < bq))
        {
          bt = ot; bq = oq; bs = os;
        }
    }
  if (gl == 0)
    {
      d.fg_ts[g] = bt; d.fg_seq[g] = bq; d.fg_slot[g] = bs;
    }
  __syncwarp (gmask);
}

/* Advance every environment: run mode until `until_ns`, gym mode until the next valid Notify().
 * SMEM: the env blob lives in shared memory during the launch. */
template <int G, bool LOG, bool GYM, bool SMEM, bool FLOW2 = false>
__global__ void __launch_bounds__ (128, LOG ? 4 : 7)
k_advance (const __grid_constant__ Static st, const LaunchArgs a)
{
  extern __shared__ __align__ (16) char smem[];
  const int lane = threadIdx.x & 31;
  const int gl = lane % G;                                   /* lane inside the group */
  const int leader = (lane / G) * G;
  const int group_in_block = threadIdx.x / G;
  const int slot_ = blockIdx.x * (blockDim.x / G) + group_in_block;
  const bool exists = slot_ < (a.env_list ? a.n_list : st.E);
  const int env = a.env_list ? (exists ? a.env_list[slot_] : 0) : slot_;
  const unsigned gmask = G == 32 ? 0xFFFFFFFFu : (((1u << G) - 1u) << leader);
  bool active = exists;
  char *gblob = a.blobs + (size_t) (active ? env : 0) * st.blob_bytes;
  char *blob = gblob;
  if (SMEM)
    {
      blob = smem + (size_t) group_in_block * st.blob_bytes;
      if (active)
        {
          const uint4 *src = (const uint4 *) gblob;
          uint4 *dst = (uint4 *) blob;
          for (int i = gl; i < st.blob_bytes / 16; i += G) dst[i] = src[i];
        }
      __syncwarp (gmask);
    }
  Dyn d;
  bind_state (st, blob, d);
  d.pool = a.pools + (size_t) (active ? env : 0) * st.P;
  d.ev_log = LOG && a.ev_logs ? a.ev_logs + (size_t) (active ? env : 0) * st.ev_cap : nullptr;
  d.st_log = LOG && a.st_logs ? a.st_logs + (size_t) (active ? env : 0) * st.st_cap : nullptr;
  Scal *sc = d.sc;
  __shared__ prb_step_rec scratch_rec[(LOG && !GYM) ? 128 / G : 1];
  prb_step_rec *rec = GYM ? &a.out[active ? env : 0] : (LOG ? &scratch_rec[group_in_block] : nullptr);
  Sim<LOG, GYM> sim (st, d, env, rec);
  const int64_t stop = st.sim_time_ns;
  const int64_t until = a.until_ns;

  int active_resume = 0;      /* gym mode: the first thing to run is the rest of the suspended Receive */
  int32_t own_action = -1;    /* in-kernel agents: the answer to the pending step */
  int steps_done = 0;
  if (GYM && G == 32 && a.qw && active)
    {
      /* the step this environment is suspended on is answered first (a.out holds its record) */
      int pend = 0;
      if (gl == 0) pend = sc->pending && !sc->finished && rec->valid;
      pend = __shfl_sync (gmask, pend, leader);
      if (pend)
        {
          int32_t act = 0;
          float tmin = INFINITY;
          if (rec->need_action)
            act = qnet_forward (st, a.qw, rec, lane, a.eps, a.seed, a.step0, st.env_id_base + (uint64_t) env, nullptr,
                                a.replay.target_mode ? &tmin : nullptr);
          if (gl == 0 && a.replay.rows) replay_record_one (st, a.replay, env, rec, act, tmin);
          __syncwarp (gmask);
          own_action = act;
        }
    }
  int fresh = 0;
  if (active && gl == 0)
    {
      if (GYM) rec->valid = 0;
      if (sc->finished)
        active = false;
      else
        {
          if (!sc->started)
            {
              sim.start ();
              fresh = 1;
            }
          if (sc->pending) active_resume = 1;
        }
    }
  active = __shfl_sync (gmask, active ? 1 : 0, leader) != 0;
  active_resume = __shfl_sync (gmask, active_resume, leader);
  fresh = __shfl_sync (gmask, fresh, leader);
  __syncwarp (gmask);

  /* flow timers sit under a two-level minimum (sim_core.cuh: scan_flow_groups): the groups are built once per episode,
   * a lane keeps the earliest of its groups in registers and looks again only after a source of one of them fired */
  constexpr bool two_level = FLOW2;        /* host: flows_two_level (st); a build of its own keeps the small-topology loop as it was */
  if (fresh && two_level)
    {
      for (int g = gl; g < st.FG; g += G) fg_rescan_serial (st, d, g);
      __syncwarp (gmask);
    }
  int64_t fts = T_IDLE; uint32_t fseq = 0xFFFFFFFFu; int fslot = -1;
  if (active)
    {
      if (two_level)
        scan_flow_groups (st, d, gl, G, fts, fseq, fslot);
      else
        scan_flows (st, d, gl, G, fts, fseq, fslot);
    }

  while (__any_sync (gmask, active))
    {
      uint32_t bk = KEY_IDLE;
      int bslot = -1;
      if (active && !active_resume)
        {
          scan_keys (st, d, gl, G, bk, bslot);
          uint32_t fk = flow_key (sc, fts);
          if (fk < bk || (fk == bk && fk != KEY_IDLE && fseq < d.t_seq[bslot]))
            {
              bk = fk;
              bslot = fslot;
            }
        }
      uint32_t mkey; int mslot;
      group_min<G> (gmask, d.t_seq, bk, bslot, mkey, mslot);
      if (!active) continue;
      int32_t action = 0;
      if (active_resume)
        {
          mslot = SLOT_RESUME;
          mkey = 0;
          action = own_action >= 0 ? own_action : (a.actions ? a.actions[env] : 0);
          own_action = -1;
          active_resume = 0;
        }
      if (mkey == KEY_FAR)
        {
          /* every armed timer is beyond the key range (rare): new epoch, exact search by the leader */
          if (gl == 0)
            {
              sim.rebase (sc->now);
              mslot = slow_min (st, d);
            }
          __syncwarp (gmask);
          mslot = __shfl_sync (gmask, mslot, leader);
        }
      int status = ST_CONTINUE;
      int go = 1;
      if (gl == 0) go = step_or_stop (st, sim, mslot, until, action, status) ? 1 : 0;
      __syncwarp (gmask);
      go = __shfl_sync (gmask, go, leader);
      if (two_level && mslot >= st.NK)
        {
          const int fg = (mslot - st.NK) >> 5;
          fg_rescan_coop<G> (st, d, fg, gl, gmask);
          if (gl == fg % G) scan_flow_groups (st, d, gl, G, fts, fseq, fslot);
        }
      else if (mslot == fslot && mslot >= 0)
        scan_flows (st, d, gl, G, fts, fseq, fslot);
      if (!go) active = false;
      if (GYM && G == 32 && a.qw)
        {
          /* the environment paused on a step: answer it here and go on (the whole warp evaluates the network) */
          status = __shfl_sync (gmask, status, leader);
          if (!go && status == ST_PAUSE && ++steps_done < a.steps)
            {
              int32_t act = 0;
              float tmin = INFINITY;
              if (rec->need_action)
                act = qnet_forward (st, a.qw, rec, lane, a.eps, a.seed, a.step0 + (uint64_t) steps_done, st.env_id_base + (uint64_t) env,
                                    nullptr, a.replay.target_mode ? &tmin : nullptr);
              if (gl == 0 && a.replay.rows) replay_record_one (st, a.replay, env, rec, act, tmin);
              __syncwarp (gmask);
              own_action = act;
              active_resume = 1;
              active = true;
            }
        }
    }
  __syncwarp (gmask);
  if (exists && gl == 0)
    {
      if (!sc->pending && until < stop && sc->now < until && !sc->finished) sc->now = until;
      sc->c.end_ns = sc->now;
    }
  if (SMEM)
    {
      __syncwarp (gmask);
      if (exists)
        {
          const uint4 *src = (const uint4 *) blob;
          uint4 *dst = (uint4 *) gblob;
          for (int i = gl; i < st.blob_bytes / 16; i += G) dst[i] = src[i];
        }
    }
}


/* ------------------------------------------------------------------------------------------------
 * Node-parallel engine (par_core.cuh): a team of T lanes per environment, lane g owns the nodes g, g+T, ...
 * One round = every lane finds the earliest pending event of its nodes and publishes its timestamp, the
 * team reduces the minimum, folds the float statistics that have become final, and every lane whose event
 * is inside the lookahead of its neighbours executes it.  The environment's state lives in shared memory
 * for the whole launch; packet records move through per-link rings in HBM.
 * ---------------------------------------------------------------------------------------------- */
struct ParArgs
{
  char *pblobs;           /* [E * pblob_bytes] */
  PRec *rings;            /* [E * ring_records] */
  prb_event *ev_logs;     /* [E * ev_cap] or null */
  int64_t until_ns;
  unsigned long long *round_stats;   /* optional: [0] rounds, [1] events executed (diagnostics) */
};

template <int T, bool LOG, bool SMEM, int CDEG = 0, int CW = 0>
#ifndef PRB_PAR_MINB
#define PRB_PAR_MINB 4          /* 128-thread blocks per SM the register allocation must allow */
#endif
__global__ void __launch_bounds__ (128, PRB_PAR_MINB)
k_run_par (const __grid_constant__ Static st, const __grid_constant__ PStatic ps, const ParArgs a)
{
  extern __shared__ __align__ (16) char smem[];
  const int lane = threadIdx.x & 31;
  const int gl = lane % T;
  const int leader = (lane / T) * T;
  const int team_in_block = threadIdx.x / T;
  const int env = blockIdx.x * (blockDim.x / T) + team_in_block;
  const unsigned tmask = T == 32 ? 0xFFFFFFFFu : (((1u << T) - 1u) << leader);
  if (env >= st.E) return;             /* whole teams leave together: every sync below names tmask only */
  char *gblob = a.pblobs + (size_t) env * ps.pblob_bytes;
  char *blob = gblob;
  if (SMEM)
    {
      blob = smem + (size_t) team_in_block * ps.pblob_bytes;
      const uint4 *src = (const uint4 *) gblob;
      uint4 *dst = (uint4 *) blob;
      for (int i = gl; i < ps.pblob_bytes / 16; i += T) dst[i] = src[i];
      __syncwarp (tmask);
    }
  PDyn d;
  p_bind_state (ps, blob, d);
  d.rings = a.rings + (size_t) env * ps.ring_records;
  d.ev_log = LOG && a.ev_logs ? a.ev_logs + (size_t) env * st.ev_cap : nullptr;
  PScal *sc = d.sc;
  PAcc acc;
  memset (&acc, 0, sizeof (acc));
  PSim<LOG, CDEG, CW> sim (st, ps, d, acc, env);
  const int N = st.N;
  const int64_t stop = st.sim_time_ns;
  const int64_t until = a.until_ns;
  const bool finished = sc->finished != 0;
  unsigned rounds = 0, executed = 0;
  bool tied = false;

  /* times inside a launch are 32-bit offsets from its start: re-base what the previous launch left */
  if (!finished && sc->base != sc->front_ts)
    {
      const uint32_t delta = (uint32_t) (sc->front_ts - sc->base);
      for (int l = gl; l < st.LA; l += T) p_rebase_link (ps, d, l, delta);
      __syncwarp (tmask);
      if (gl == 0) sc->base = sc->front_ts;
      __syncwarp (tmask);
      sim.set_window (until);
      for (int x = gl; x < N; x += T) sim.rescan_flows (x);
    }
  sim.set_window (until);
  if (!finished && !sc->started)
    for (int x = gl; x < N; x += T) sim.start_node (x);
  __syncwarp (tmask);

  /* every statistics record older than `bound`, in timestamp order; one lane applies at a time */
  auto fold = [&] (uint32_t bound) {
    for (;;)
      {
        uint32_t ft = P_FAR32;
        int fx = -1;
        for (int x = gl; x < N; x += T)
          if (d.s_head[x] != d.s_tail[x])
            {
              uint32_t ts = d.sring[x * P_SR + (d.s_head[x] & (P_SR - 1))].ts;
              if (ts < bound && ts < ft)
                {
                  ft = ts; fx = x;
                }
            }
        uint32_t m = __reduce_min_sync (tmask, ft);
        if (m == P_FAR32) break;
        unsigned b = __ballot_sync (tmask, ft == m);
        if (b & (b - 1))
          {
            /* one timestamp on several lanes: the record whose event was scheduled earliest */
            uint32_t inv = ft == m ? 0x3FFFFFFFu - (d.sring[fx * P_SR + (d.s_head[fx] & (P_SR - 1))].code & 0x3FFFFFFFu) : P_FAR32;
            uint32_t mi = __reduce_min_sync (tmask, inv);
            b = __ballot_sync (tmask, inv == mi);
          }
        if (lane == __ffs (b) - 1)
          {
            p_apply_stat (sc->c, d.sring[fx * P_SR + (d.s_head[fx] & (P_SR - 1))]);
            d.s_head[fx] = (uint8_t) (d.s_head[fx] + 1);
          }
        __syncwarp (tmask);
      }
  };

  bool stats_pending = false, stats_tight = false;      /* team-uniform: some ring holds a record / is more than half full */
  while (!finished)
    {
      /* phase 1: earliest pending event of my nodes */
      uint32_t mt = P_FAR32;
      bool mine_in_time = false;
      for (int x = gl; x < N; x += T)
        {
          PCand c = sim.select (x);
          sim.prefetch_record (c);
          d.t_next[x] = c.ts;
          d.cand[x] = (c.kind << 24) | (c.idx & 0xFFFFFF);
          mt = c.ts < mt ? c.ts : mt;
          mine_in_time = mine_in_time || sim.in_time (c);
        }
      __syncwarp (tmask);
      const bool more = __any_sync (tmask, mine_in_time);
      const uint32_t gmin = __reduce_min_sync (tmask, mt);
      if (__any_sync (tmask, (acc.errors & (PRB_ERR_TIE | PRB_ERR_POOL)) != 0))
        {
          /* the order of two simultaneous events could not be rebuilt, or a link ring is full (sized for the
           * usual mix of packet sizes, not for a queue of nothing but the smallest ones): this environment is
           * handed over to the event-serial engine (host: resolve_ties), which replays it from its reset state */
          tied = true;
          break;
        }
#ifndef PRB_FOLD_EVERY
#define PRB_FOLD_EVERY 16
#endif
      /* the rings hold 8 records and an event pushes at most 2: folded every 16th round, at once when some ring is more than
       * half full (a node with less than two free records waits, see safe), and at the end of the run everything that is left */
      if ((stats_pending && (stats_tight || rounds % PRB_FOLD_EVERY == 0)) || !more) fold (more ? gmin : P_FAR32);
      if (!more) break;
      /* phase 2: the earliest of my events that nothing can precede any more */
      PCand my;
      my.ts = P_FAR32; my.kind = PK_NONE; my.idx = -1;
      int myx = -1;
      for (int x = gl; x < N; x += T)
        {
          PCand c;
          c.ts = d.t_next[x]; c.kind = d.cand[x] >> 24; c.idx = d.cand[x] & 0xFFFFFF;
          if (c.ts < my.ts && sim.safe (x, c, gmin))
            {
              my = c; myx = x;
            }
        }
      unsigned go = __ballot_sync (tmask, myx >= 0);     /* also orders the reads above before the handlers */
      if (go == 0)
        {
          acc.errors |= PRB_ERR_SEQ;                       /* cannot happen: the node that holds gmin is safe */
          break;
        }
      if (myx >= 0) sim.execute (myx, my);
      __syncwarp (tmask);
      unsigned left = 0;
      for (int x = gl; x < N; x += T)
        {
          const unsigned n = (uint8_t) (d.s_tail[x] - d.s_head[x]);
          left |= (n != 0 ? 1u : 0u) | (n > (unsigned) P_SR / 2 ? 2u : 0u);
        }
      left = __reduce_or_sync (tmask, left);
      stats_pending = (left & 1u) != 0;
      stats_tight = (left & 2u) != 0;
      rounds++;
      executed += __popc (go);
    }
  __syncwarp (tmask);
  if (!finished)
    {
      p_flush_acc (sc, acc);
      __syncwarp (tmask);
      if (gl == 0)
        {
          sc->started = 1;
          sc->front_ts = until < stop ? until : stop;
          if (until >= stop) sc->finished = 1;
          if (tied)
            {
              sc->finished = 1;
              sc->c.errors |= PRB_ERR_TIE;       /* marks the hand-over, whatever caused it */
              if (a.round_stats) atomicAdd (&a.round_stats[2], 1ull);
            }
          if (a.round_stats)
            {
              atomicAdd (&a.round_stats[0], (unsigned long long) rounds);
              atomicAdd (&a.round_stats[1], (unsigned long long) executed);
            }
        }
    }
  if (SMEM)
    {
      __syncwarp (tmask);
      const uint4 *src = (const uint4 *) blob;
      uint4 *dst = (uint4 *) gblob;
      for (int i = gl; i < ps.pblob_bytes / 16; i += T) dst[i] = src[i];
    }
}

__global__ void
k_reset_par (const __grid_constant__ Static st, const __grid_constant__ PStatic ps, char *pblobs, const char *pblob0)
{
  const int64_t words = ps.pblob_bytes / 16;
  const int64_t total = words * st.E;
  const int64_t stride = (int64_t) gridDim.x * blockDim.x;
  for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride)
    ((uint4 *) pblobs)[i] = ((const uint4 *) pblob0)[i % words];
}

__global__ void
k_counters_par (const __grid_constant__ Static st, const __grid_constant__ PStatic ps, const char *pblobs, prb_counters *out)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= st.E) return;
  PDyn d;
  p_bind_state (ps, const_cast<char *> (pblobs) + (size_t) e * ps.pblob_bytes, d);
  prb_counters c;
  p_finalize_counters (st, d, c);
  out[e] = c;
}

__global__ void
k_queue_bytes_par (const __grid_constant__ Static st, const __grid_constant__ PStatic ps, const char *pblobs, int e, int32_t *out)
{
  int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= st.L) return;
  const uint32_t *q = (const uint32_t *) (pblobs + (size_t) e * ps.pblob_bytes + ps.off_q_bytes);
  out[l] = (int32_t) q[l];
}

__global__ void
k_log_counts_par (const __grid_constant__ PStatic ps, const char *pblobs, int e, int64_t *out)
{
  const PScal *sc = (const PScal *) (pblobs + (size_t) e * ps.pblob_bytes + ps.off_scal);
  out[0] = sc->ev_n;
  out[1] = 0;
}


/* reset: every env gets the initial blob, its policy RNG, and a fully linked free list */
__global__ void
k_reset (const __grid_constant__ Static st, char *blobs, Pkt *pools, const char *blob0)
{
  const int64_t words = st.blob_bytes / 16;
  const int64_t total = words * st.E;
  const int64_t stride = (int64_t) gridDim.x * blockDim.x;
  const int64_t tid = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  for (int64_t i = tid; i < total; i += stride)
    ((uint4 *) blobs)[i] = ((const uint4 *) blob0)[i % words];
  const int64_t np = (int64_t) st.P * st.E;
  for (int64_t i = tid; i < np; i += stride)
    {
      int k = (int) (i % st.P);
      pools[i].next = (uint16_t) (k + 1 < st.P ? k + 1 : NIL);
    }
}

__global__ void
k_seed_policy (const __grid_constant__ Static st, char *blobs)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= st.E) return;
  Scal *sc = (Scal *) (blobs + (size_t) e * st.blob_bytes + st.off_scal);
  uint64_t r = (st.policy_seed + (uint64_t) e) * 0x9E3779B97F4A7C15ULL + 0x1234567ULL;
  sc->pol_rng = r ? r : 1;
  sc->pol_pos = 0;
}

__global__ void
k_tied_flags (const __grid_constant__ Static st, const __grid_constant__ PStatic ps, const char *pblobs, uint8_t *flags)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= st.E) return;
  PDyn d;
  p_bind_state (ps, const_cast<char *> (pblobs) + (size_t) e * ps.pblob_bytes, d);
  flags[e] = (d.sc->c.errors & PRB_ERR_TIE) ? 1 : 0;
}

__global__ void
k_counters_list (const __grid_constant__ Static st, const char *blobs, const int32_t *list, int n, prb_counters *out)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int e = list[i];
  Dyn d;
  bind_state (st, const_cast<char *> (blobs) + (size_t) e * st.blob_bytes, d);
  prb_counters c;
  finalize_counters (st, d, c);
  out[e] = c;
}

__global__ void
k_counters (const __grid_constant__ Static st, const char *blobs, prb_counters *out)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= st.E) return;
  Dyn d;
  bind_state (st, const_cast<char *> (blobs) + (size_t) e * st.blob_bytes, d);
  prb_counters c;
  finalize_counters (st, d, c);
  out[e] = c;
}

__global__ void
k_queue_bytes (const __grid_constant__ Static st, const char *blobs, int e, int32_t *out)
{
  int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= st.L) return;
  const uint32_t *q = (const uint32_t *) (blobs + (size_t) e * st.blob_bytes + st.off_q_bytes);
  out[l] = (int32_t) q[l];
}

__global__ void
k_log_counts (const __grid_constant__ Static st, const char *blobs, int e, int64_t *out)
{
  const Scal *sc = (const Scal *) (blobs + (size_t) e * st.blob_bytes + st.off_scal);
  out[0] = sc->ev_n;
  out[1] = sc->st_n;
}

/* ------------------------------------------------------------------------------------------------
 * Host side
 * ---------------------------------------------------------------------------------------------- */

struct prb_env
{
  HostModel hm;
  Static dst;                 /* Static with DEVICE table pointers */
  HostTrace tr;
  int device = 0;
  cudaStream_t stream = nullptr;
  int G = 32;
  int max_smem = 0, smem_per_sm = 0, num_sms = 148;
  char *d_tables = nullptr;   /* one allocation for all read-only tables */
  char *d_blob0 = nullptr;
  char *d_blobs = nullptr;
  Pkt *d_pools = nullptr;
  prb_event *d_ev = nullptr;
  prb_step_rec *d_st = nullptr;
  int32_t *d_pol = nullptr;   int64_t pol_cap = 0;
  char *d_trace = nullptr;
  int32_t *d_flow_first = nullptr;
  prb_counters *d_counters = nullptr;
  prb_step_rec *d_out = nullptr;
  int32_t *d_actions = nullptr;
  int64_t *d_scratch = nullptr;
  int64_t launches = 0;
  bool log_on = false;
  /* node-parallel engine */
  PModel pm;
  PStatic dps;                /* PStatic with DEVICE table pointers */
  char *d_ptables = nullptr;
  char *d_pblob0 = nullptr;
  char *d_pblobs = nullptr;   /* allocated on first use */
  PRec *d_rings = nullptr;
  int32_t *d_tr_prev = nullptr;
  unsigned long long *d_round_stats = nullptr;
  int engine_mode = PRB_ENGINE_AUTO;
  int engine_active = 0;      /* engine that has advanced the environments since the last reset (0: none) */
  int team = 0;
  int64_t par_front_ns = 0;
  bool pol_node_trace = false;
  /* environments the parallel engine handed over to the serial engine (undecided order of simultaneous events) */
  std::vector<int32_t> tied;
  std::vector<uint8_t> tied_flag;
  int32_t *d_tied_list = nullptr;
  uint8_t *d_tied_flags = nullptr;
  int64_t par_until_arg = 0;       /* until_ns of the last parallel launch */
  int64_t tied_until = 0;          /* until_ns the handed-over environments have been advanced to */
  /* Table agents installed during this episode, by the simulated time from which the parallel launches used them:
   * a replay by the serial engine walks the same segments (prb_set_policy between prb_run calls is legal). */
  struct PolSeg { int64_t from_ns; std::vector<int32_t> table; };
  std::vector<PolSeg> pol_segs;
  std::vector<int32_t> pol_host;   /* host copy of the table policy in force */
  int32_t *d_pol_replay = nullptr;
  double replay_ms = 0.0;          /* device time of the serial replays of this episode */
  cudaEvent_t rp_ev0 = nullptr, rp_ev1 = nullptr;
  /* dqn_buffer agents */
  float *d_qw = nullptr;
  uint64_t q_step = 0;
  bool qnet_unfused = false;   /* PRB_QNET_UNFUSED=1: prb_step_qnet as a sequence of launches (A/B tests) */
  int qnet_mode = PRB_QNET_FUSED;  /* prb_qnet_set_mode */
  int32_t *d_qcount = nullptr, *d_qperm = nullptr;
  int q_parity = 0;
  /* learner on the device (prb_learner_*): the online weights ARE d_qw */
  float *d_lw_target = nullptr, *d_lexp = nullptr, *d_lexp_sq = nullptr, *d_lsteps = nullptr, *d_ltd = nullptr;
  float l_lr = 1e-3f, l_gamma = 1.0f, l_b1 = 0.9f, l_b2 = 0.999f, l_eps = 1e-7f;
  int l_target_mode = 0;
  /* replay buffers */
  ReplayArgs rp = {};
  uint64_t rp_call = 0;
};

template <typename T>
static const T *
place (std::vector<char> &buf, const std::vector<T> &v, size_t &off_out)
{
  size_t off = (buf.size () + 15) / 16 * 16;
  buf.resize (off + std::max<size_t> (v.size () * sizeof (T), 16));
  if (!v.empty ()) memcpy (buf.data () + off, v.data (), v.size () * sizeof (T));
  off_out = off;
  return nullptr;
}

static int
upload_tables (prb_env *env)
{
  HostModel &hm = env->hm;
  std::vector<char> buf;
  size_t o[16];
  place (buf, hm.link_from, o[0]); place (buf, hm.link_to, o[1]); place (buf, hm.link_rc, o[2]);
  place (buf, hm.tx_ns, o[3]); place (buf, hm.szc_bytes, o[4]); place (buf, hm.route, o[5]);
  place (buf, hm.map_overlay, o[6]); place (buf, hm.o2u, o[7]); place (buf, hm.agent_k, o[8]);
  place (buf, hm.agent_nbr, o[9]); place (buf, hm.agent_sign_szc, o[10]); place (buf, hm.flow_src, o[11]);
  place (buf, hm.flow_dst, o[12]); place (buf, hm.flow_mean_s, o[13]); place (buf, hm.opt_rejected, o[14]);
  CUDA_TRY (cudaMalloc (&env->d_tables, buf.size ()));
  CUDA_TRY (cudaMemcpyAsync (env->d_tables, buf.data (), buf.size (), cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  Static &s = env->dst;
  s = hm.st;
  char *b = env->d_tables;
  s.link_from = (const int16_t *) (b + o[0]); s.link_to = (const int16_t *) (b + o[1]);
  s.link_rc = (const uint8_t *) (b + o[2]); s.tx_ns = (const int64_t *) (b + o[3]);
  s.szc_bytes = (const uint16_t *) (b + o[4]); s.route = (const int16_t *) (b + o[5]);
  s.map_overlay = (const int16_t *) (b + o[6]); s.o2u = (const int16_t *) (b + o[7]);
  s.agent_k = (const uint8_t *) (b + o[8]); s.agent_nbr = (const int16_t *) (b + o[9]);
  s.agent_sign_szc = (const uint8_t *) (b + o[10]); s.flow_src = (const int16_t *) (b + o[11]);
  s.flow_dst = (const int16_t *) (b + o[12]); s.flow_mean_s = (const double *) (b + o[13]);
  s.opt_rejected = hm.opt_rejected.empty () ? nullptr : (const double *) (b + o[14]);
  s.pol_data = nullptr; s.pol_n = 0;
  s.tr_t = nullptr; s.tr_uid = nullptr; s.tr_flags = nullptr; s.tr_next = nullptr; s.flow_first = nullptr;
  return PRB_OK;
}

static int
upload_ptables (prb_env *env)
{
  PModel &pm = env->pm;
  env->dps = pm.ps;
  if (!pm.supported) return PRB_OK;
  std::vector<char> buf;
  size_t o[8];
  place (buf, pm.in_links, o[0]); place (buf, pm.out_links, o[1]); place (buf, pm.ring_off, o[2]);
  place (buf, pm.ring_mask, o[3]); place (buf, pm.flow_begin, o[4]);
  CUDA_TRY (cudaMalloc (&env->d_ptables, buf.size ()));
  CUDA_TRY (cudaMemcpyAsync (env->d_ptables, buf.data (), buf.size (), cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaMalloc (&env->d_pblob0, pm.ps.pblob_bytes));
  CUDA_TRY (cudaMemcpyAsync (env->d_pblob0, pm.pblob0.data (), pm.ps.pblob_bytes, cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaMalloc (&env->d_round_stats, 32));
  CUDA_TRY (cudaMemsetAsync (env->d_round_stats, 0, 32, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  PStatic &s = env->dps;
  char *b = env->d_ptables;
  s.in_links = (const int16_t *) (b + o[0]); s.out_links = (const int16_t *) (b + o[1]);
  s.ring_off = (const int32_t *) (b + o[2]); s.ring_mask = (const uint16_t *) (b + o[3]);
  s.flow_begin = (const int32_t *) (b + o[4]);
  s.tr_prev = nullptr; s.pol_off = nullptr;
  return PRB_OK;
}

/* the parallel engine's state is allocated when it is first used */
static int
ensure_par_state (prb_env *env)
{
  if (env->d_pblobs) return PRB_OK;
  const Static &st = env->dst;
  const PStatic &ps = env->dps;
  cudaError_t e = cudaMalloc (&env->d_pblobs, (size_t) ps.pblob_bytes * st.E);
  if (e == cudaSuccess) e = cudaMalloc (&env->d_rings, sizeof (PRec) * (size_t) ps.ring_records * st.E);
  if (e != cudaSuccess) return fail (PRB_ENOMEM, std::string ("cudaMalloc (parallel engine state): ") + cudaGetErrorString (e));
  /* no record has been written yet: every lineage field reads as "no transmission start" (P_XW_QUEUED) */
  CUDA_TRY (cudaMemsetAsync (env->d_rings, 0xFF, sizeof (PRec) * (size_t) ps.ring_records * st.E, env->stream));
  int64_t work = (int64_t) ps.pblob_bytes / 16 * st.E;
  int blocks = (int) std::min<int64_t> ((work + 255) / 256, 148 * 16);
  k_reset_par<<<blocks, 256, 0, env->stream>>> (st, ps, env->d_pblobs, env->d_pblob0);
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  return PRB_OK;
}

template <int T, bool LOG>
static int
launch_par_t (prb_env *env, const ParArgs &a)
{
  const Static &st = env->dst;
  const PStatic &ps = env->dps;
  /* block size: the one that keeps the most environments resident per SM (shared memory holds the state
   * of every team of a block; 1 KB per block is reserved by the system); among equals the one that leaves the
   * fewest environments on the busiest SM — the step ends with the slowest SM, and warps of one SM slow each
   * other down through the instruction cache (4096 Abilene environments: 128-thread blocks put 32 on 68 of the
   * 148 SMs, 64-thread blocks at most 28: 12.49 -> 12.28 ms) */
  int threads = 0, best = -1, best_peak = 0;
  const char *forced = getenv ("PRB_PAR_THREADS");          /* experiments: block size of k_run_par */
  for (int t = 128; t >= 32 && t >= T; t >>= 1)
    {
      if (forced && atoi (forced) != t) continue;
      int teams = t / T;
      size_t per_block = (size_t) teams * ps.pblob_bytes + 1024;
      if (per_block > (size_t) env->max_smem) continue;
      int blocks_sm = (int) std::min<size_t> ((size_t) env->smem_per_sm / per_block, 32);
      blocks_sm = std::min (blocks_sm, PRB_PAR_MINB * 128 / t);  /* registers: __launch_bounds__ of k_run_par */
      int resident = blocks_sm * teams;
      int blocks_total = (st.E + teams - 1) / teams;
      int peak = std::min ((blocks_total + env->num_sms - 1) / std::max (env->num_sms, 1), blocks_sm) * teams;
      if (resident > best || (resident == best && peak < best_peak))
        {
          best = resident; threads = t; best_peak = peak;
        }
    }
  if (threads > 0)
    {
      const int teams = threads / T;
      const int blocks = (st.E + teams - 1) / teams;
      size_t smem = (size_t) teams * ps.pblob_bytes;
      /* the common shape (row length 3, 8 wire slots: Abilene-like degrees at 1 ms / 500 kbit/s) has its own build of
       * the throughput kernel with both as immediates */
      auto k = k_run_par<T, LOG, true>;
      if (T == 16 && !LOG && ps.DEG == 3 && ps.W == 8) k = k_run_par<T, LOG, true, (T == 16 && !LOG) ? 3 : 0, (T == 16 && !LOG) ? 8 : 0>;
      CUDA_TRY (cudaFuncSetAttribute (k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
      k<<<blocks, threads, smem, env->stream>>> (st, ps, a);
    }
  else
    {
      const int teams = 128 / T;
      k_run_par<T, LOG, false><<<(st.E + teams - 1) / teams, 128, 0, env->stream>>> (st, ps, a);
    }
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  return PRB_OK;
}

static int
launch_par (prb_env *env, const ParArgs &a)
{
  bool log = env->log_on;
  switch (env->team)
    {
    case 8: return log ? launch_par_t<8, true> (env, a) : launch_par_t<8, false> (env, a);
    case 16: return log ? launch_par_t<16, true> (env, a) : launch_par_t<16, false> (env, a);
    case 32: return log ? launch_par_t<32, true> (env, a) : launch_par_t<32, false> (env, a);
    default: return fail (PRB_EINVAL, "team size must be 8, 16 or 32");
    }
}

/* prb_run on the parallel engine; launches cover at most 4 simulated seconds each (32-bit time offsets) */
static int
run_parallel (prb_env *env, int64_t until_ns)
{
  int rc = ensure_par_state (env);
  if (rc != PRB_OK) return rc;
  const int64_t stop = env->dst.sim_time_ns;
  int64_t target = std::min (until_ns, stop);
  /* nothing lies before the frontier: times inside a launch are unsigned offsets from it and the frontier never
   * moves backwards (the serial engine treats such a call as a no-op too) */
  if (env->engine_active == PRB_ENGINE_PARALLEL && target <= env->par_front_ns) return PRB_OK;
  if (env->dst.policy == PRB_POLICY_TABLE && (env->pol_segs.empty () || env->pol_segs.back ().table != env->pol_host))
    {
      prb_env::PolSeg s;
      s.from_ns = env->engine_active == PRB_ENGINE_PARALLEL ? env->par_front_ns : 0;
      s.table = env->pol_host;
      if (!env->pol_segs.empty () && env->pol_segs.back ().from_ns == s.from_ns) env->pol_segs.pop_back ();
      env->pol_segs.push_back (std::move (s));
    }
  ParArgs a;
  a.pblobs = env->d_pblobs; a.rings = env->d_rings; a.ev_logs = env->d_ev; a.round_stats = env->d_round_stats;
  do
    {
      int64_t chunk_end = std::min<int64_t> (target, env->par_front_ns + env->pm.ps.max_chunk_ns);
      a.until_ns = chunk_end >= stop ? std::max<int64_t> (until_ns, stop) : chunk_end;
      if ((rc = launch_par (env, a)) != PRB_OK) return rc;
      env->par_until_arg = a.until_ns;
      env->par_front_ns = chunk_end;
    }
  while (env->par_front_ns < target);
  env->engine_active = PRB_ENGINE_PARALLEL;
  return PRB_OK;
}

template <int G, bool LOG, bool GYM>
static int
launch_advance_g (prb_env *env, const LaunchArgs &a)
{
  const Static &st = env->dst;
  const int threads = 128;
  const int envs_per_block = threads / G;
  const int n_envs = a.env_list ? a.n_list : st.E;
  const int blocks = (n_envs + envs_per_block - 1) / envs_per_block;
  size_t smem = (size_t) envs_per_block * st.blob_bytes;
  /* shared-memory staging whenever the blobs of a block fit: it pays even for a single Gym step (a launch lasts as long
   * as the environment that needs the most events to reach its next decision: 137 -> 114 us at 4096 Abilene environments) */
  bool use_smem = smem + sizeof (prb_step_rec) * envs_per_block + 1024 <= (size_t) env->max_smem;
  if (use_smem)
    {
      auto k = flows_two_level (st) ? k_advance<G, LOG, GYM, true, true> : k_advance<G, LOG, GYM, true, false>;
      CUDA_TRY (cudaFuncSetAttribute (k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
      k<<<blocks, threads, smem, env->stream>>> (st, a);
    }
  else
    {
      if (flows_two_level (st))
        k_advance<G, LOG, GYM, false, true><<<blocks, threads, 0, env->stream>>> (st, a);
      else
        k_advance<G, LOG, GYM, false, false><<<blocks, threads, 0, env->stream>>> (st, a);
    }
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  return PRB_OK;
}

static int
launch_advance (prb_env *env, const LaunchArgs &a)
{
  bool log = env->log_on;
  bool gym = env->dst.mode == MODE_GYM;
#define PRB_DISPATCH(G_)                                                                              \
  if (gym) return log ? launch_advance_g<G_, true, true> (env, a) : launch_advance_g<G_, false, true> (env, a); \
  return log ? launch_advance_g<G_, true, false> (env, a) : launch_advance_g<G_, false, false> (env, a);
  switch (env->G)
    {
    case 32: PRB_DISPATCH (32)
    case 16: PRB_DISPATCH (16)
    case 8: PRB_DISPATCH (8)
    default: return fail (PRB_EINVAL, "lanes_per_env must be 8, 16 or 32");
    }
#undef PRB_DISPATCH
}

/* ---- prb_queue_bytes without any allocation: the scratch is a file-scope device array ----
 * g_noalloc_scratch is a HOST-side stand-in in host code, not a device address: its device address is looked up once
 * per device (noalloc_setup, from prb_create with that device current) and that pointer is what the kernels get. */
static __device__ __align__ (256) int32_t g_noalloc_scratch[PRB_NOALLOC_MAX_LINKS];
static int32_t *g_noalloc_ptr[PRB_NOALLOC_MAX_DEVICES];

static int
noalloc_setup (int device)
{
  if (device < 0 || device >= PRB_NOALLOC_MAX_DEVICES) return PRB_OK;     /* such a device gets PRB_ESTATE from the entry point */
  if (g_noalloc_ptr[device]) return PRB_OK;
  int32_t *p = nullptr;
  CUDA_TRY (cudaGetSymbolAddress ((void **) &p, g_noalloc_scratch));
  g_noalloc_ptr[device] = p;
  return PRB_OK;
}

extern "C" {

const char *prb_last_error (void) { return g_err.c_str (); }
const char *prb_version (void) { return PRB_VERSION; }
int64_t prb_seconds_to_ns (double s) { return host_seconds_to_ns (s); }

double
prb_ns_to_seconds (int64_t ns)
{
  Static st;
  host_invert_1e9 (&st.inv1e9_hi, &st.inv1e9_lo);
  return ns < 0 ? -ns_to_seconds (st, -ns) : ns_to_seconds (st, ns);
}

int64_t prb_tx_time_ns (uint32_t bytes, uint64_t bps) { return host_tx_time_ns (bytes, bps); }
int32_t prb_build_links (int32_t n, const uint8_t *adj, int32_t *from, int32_t *to) { return host_build_links (n, adj, from, to); }
int32_t prb_global_routes (int32_t n, const uint8_t *adj, int32_t *route_out) { return host_global_routes (n, adj, route_out); }

int
prb_destroy (prb_env *env)
{
  if (!env) return PRB_OK;
  cudaSetDevice (env->device);
  cudaStreamSynchronize (env->stream);
  cudaFree (env->d_tables); cudaFree (env->d_blob0); cudaFree (env->d_blobs); cudaFree (env->d_pools);
  cudaFree (env->d_ev); cudaFree (env->d_st); cudaFree (env->d_pol); cudaFree (env->d_trace);
  cudaFree (env->d_flow_first); cudaFree (env->d_counters); cudaFree (env->d_out); cudaFree (env->d_actions);
  cudaFree (env->d_scratch);
  cudaFree (env->d_ptables); cudaFree (env->d_pblob0); cudaFree (env->d_pblobs); cudaFree (env->d_rings);
  cudaFree (env->d_tr_prev); cudaFree (env->d_round_stats); cudaFree (env->d_qw); cudaFree (env->d_tied_list); cudaFree (env->d_tied_flags); cudaFree (env->d_pol_replay);
  cudaFree (env->d_qcount); cudaFree (env->d_qperm);
  cudaFree (env->d_lw_target); cudaFree (env->d_lexp); cudaFree (env->d_lexp_sq); cudaFree (env->d_lsteps); cudaFree (env->d_ltd);
  if (env->rp_ev0) { cudaEventDestroy (env->rp_ev0); cudaEventDestroy (env->rp_ev1); }
  cudaFree (env->rp.rows); cudaFree (env->rp.cursor); cudaFree (env->rp.pending); cudaFree (env->rp.tmin);
  cudaFree (env->rp.row_gstep); cudaFree (env->rp.latest); cudaFree (env->rp.psum);
  delete env;
  return PRB_OK;
}

int
prb_create (const prb_config *cfg, int device, void *cuda_stream, prb_env **out)
{
  if (!cfg || !out) return fail (PRB_EINVAL, "null argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount (&ndev) != cudaSuccess || ndev <= 0)
    return fail (PRB_ECUDA, "no CUDA device: prisma_b200 has no CPU execution path");
  if (device < 0 || device >= ndev) return fail (PRB_EINVAL, "bad device index");
  prb_env *env = new prb_env ();
  if (!build_model (*cfg, env->hm))
    {
      std::string msg = env->hm.error;
      delete env;
      return fail (PRB_EINVAL, msg);
    }
  env->device = device;
  env->stream = (cudaStream_t) cuda_stream;
  env->G = cfg->lanes_per_env > 0 ? cfg->lanes_per_env : 32;
  env->qnet_unfused = getenv ("PRB_QNET_UNFUSED") != nullptr;
  if (const char *qm = getenv ("PRB_QNET_MODE"))
    env->qnet_mode = !strcmp (qm, "mma") ? PRB_QNET_MMA : (!strcmp (qm, "warp") ? PRB_QNET_WARP : PRB_QNET_FUSED);
  if (env->G != 32 && env->G != 16 && env->G != 8)
    {
      delete env;
      return fail (PRB_EINVAL, "lanes_per_env must be 8, 16 or 32");
    }
  int rc = PRB_OK;
  auto guard = [&] (int r) { if (r != PRB_OK) { prb_destroy (env); } return r; };
  if (cudaSetDevice (device) != cudaSuccess) return guard (fail (PRB_ECUDA, "cudaSetDevice failed"));
  if ((rc = noalloc_setup (device)) != PRB_OK) return guard (rc);       /* prb_queue_bytes_noalloc: address of the static scratch */
  cudaDeviceGetAttribute (&env->max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
  cudaDeviceGetAttribute (&env->smem_per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, device);
  cudaDeviceGetAttribute (&env->num_sms, cudaDevAttrMultiProcessorCount, device);
  if ((rc = upload_tables (env)) != PRB_OK) return guard (rc);
  build_pmodel (*cfg, env->hm, env->pm);      /* pm.supported says whether prb_run may use the parallel engine */
  env->team = env->hm.st.N <= 8 ? 8 : (env->hm.st.N <= 16 ? 16 : 32);
  if ((rc = upload_ptables (env)) != PRB_OK) return guard (rc);
  const Static &st = env->dst;
  auto alloc = [&] (void **p, size_t bytes) -> int {
    cudaError_t e = cudaMalloc (p, bytes);
    if (e != cudaSuccess) return fail (PRB_ENOMEM, std::string ("cudaMalloc: ") + cudaGetErrorString (e));
    return PRB_OK;
  };
  if ((rc = alloc ((void **) &env->d_blob0, st.blob_bytes)) != PRB_OK) return guard (rc);
  if ((rc = alloc ((void **) &env->d_blobs, (size_t) st.blob_bytes * st.E)) != PRB_OK) return guard (rc);
  if ((rc = alloc ((void **) &env->d_pools, sizeof (Pkt) * (size_t) st.P * st.E)) != PRB_OK) return guard (rc);
  if ((rc = alloc ((void **) &env->d_counters, sizeof (prb_counters) * (size_t) st.E)) != PRB_OK) return guard (rc);
  if ((rc = alloc ((void **) &env->d_out, sizeof (prb_step_rec) * (size_t) st.E)) != PRB_OK) return guard (rc);
  if ((rc = alloc ((void **) &env->d_actions, sizeof (int32_t) * (size_t) st.E)) != PRB_OK) return guard (rc);
  if ((rc = alloc ((void **) &env->d_scratch, 4096)) != PRB_OK) return guard (rc);
  if ((rc = alloc ((void **) &env->d_flow_first, sizeof (int32_t) * std::max (st.F, 1))) != PRB_OK) return guard (rc);
  if (cudaMemcpyAsync (env->d_blob0, env->hm.blob0.data (), st.blob_bytes, cudaMemcpyHostToDevice, env->stream) != cudaSuccess)
    return guard (fail (PRB_ECUDA, "upload of the initial state failed"));
  if (cudaMemsetAsync (env->d_actions, 0, sizeof (int32_t) * (size_t) st.E, env->stream) != cudaSuccess)
    return guard (fail (PRB_ECUDA, "memset failed"));
  *out = env;
  rc = prb_reset (env);
  if (rc != PRB_OK)
    {
      *out = nullptr;
      return guard (rc);
    }
  return PRB_OK;
}

int
prb_reset (prb_env *env)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  int64_t work = std::max<int64_t> ((int64_t) st.blob_bytes / 16 * st.E, (int64_t) st.P * st.E);
  int blocks = (int) std::min<int64_t> ((work + 255) / 256, 148 * 16);
  k_reset<<<blocks, 256, 0, env->stream>>> (st, env->d_blobs, env->d_pools, env->d_blob0);
  CUDA_TRY (cudaGetLastError ());
  k_seed_policy<<<(st.E + 127) / 128, 128, 0, env->stream>>> (st, env->d_blobs);
  CUDA_TRY (cudaGetLastError ());
  env->launches += 2;
  if (env->d_pblobs)
    {
      int64_t pwork = (int64_t) env->dps.pblob_bytes / 16 * st.E;
      int pblocks = (int) std::min<int64_t> ((pwork + 255) / 256, 148 * 16);
      k_reset_par<<<pblocks, 256, 0, env->stream>>> (st, env->dps, env->d_pblobs, env->d_pblob0);
      CUDA_TRY (cudaGetLastError ());
      env->launches++;
    }
  if (env->rp.pending)   /* Agent.temp_obs belongs to the episode; the stored transitions survive it */
    CUDA_TRY (cudaMemsetAsync (env->rp.pending, 0xFF, sizeof (int32_t) * (size_t) st.E * (size_t) env->rp.pcap * env->rp.PW, env->stream));
  env->engine_active = 0;
  env->par_front_ns = 0;
  env->tied.clear ();
  std::fill (env->tied_flag.begin (), env->tied_flag.end (), 0);
  env->tied_until = 0;
  env->pol_segs.clear ();
  env->replay_ms = 0.0;
  if (env->d_round_stats) CUDA_TRY (cudaMemsetAsync (env->d_round_stats + 2, 0, 8, env->stream));
  return PRB_OK;
}

/* The serial engine advances the first n environments of d_tied_list from `from_ns` to `to_ns`, stretch by stretch
 * with the table agent the parallel launches ran that stretch of simulated time with (pol_segs); the device time of
 * these launches is accumulated in replay_ms (prb_handover_stats). */
static int
replay_span (prb_env *env, int32_t n, int64_t from_ns, int64_t to_ns)
{
  if (!env->rp_ev0)
    {
      CUDA_TRY (cudaEventCreate (&env->rp_ev0));
      CUDA_TRY (cudaEventCreate (&env->rp_ev1));
    }
  CUDA_TRY (cudaEventRecord (env->rp_ev0, env->stream));
  const size_t nseg = env->pol_segs.size ();
  const int32_t *pol_saved = env->dst.pol_data;
  for (size_t i = 0; i < std::max<size_t> (nseg, 1); i++)
    {
      int64_t seg_to = i + 1 < nseg ? std::min (env->pol_segs[i + 1].from_ns, to_ns) : to_ns;
      if (seg_to <= from_ns) continue;
      LaunchArgs a;
      a.blobs = env->d_blobs; a.pools = env->d_pools; a.ev_logs = nullptr; a.st_logs = nullptr;
      a.actions = nullptr; a.out = nullptr; a.until_ns = seg_to;
      a.qw = nullptr; a.steps = 1; a.eps = 0.0f; a.seed = 0; a.step0 = 0; a.replay = ReplayArgs ();
      a.env_list = env->d_tied_list; a.n_list = n;
      if (nseg > 1)
        {
          const std::vector<int32_t> &t = env->pol_segs[i].table;
          if (!env->d_pol_replay) CUDA_TRY (cudaMalloc (&env->d_pol_replay, sizeof (int32_t) * t.size ()));
          CUDA_TRY (cudaMemcpyAsync (env->d_pol_replay, t.data (), sizeof (int32_t) * t.size (), cudaMemcpyHostToDevice, env->stream));
          env->dst.pol_data = env->d_pol_replay;
        }
      env->dst.mode = MODE_RUN;
      int rc = launch_advance (env, a);
      env->dst.pol_data = pol_saved;
      if (rc != PRB_OK) return rc;
      if (nseg > 1) CUDA_TRY (cudaStreamSynchronize (env->stream));    /* d_pol_replay is reused by the next stretch */
      if (seg_to >= to_ns) break;
    }
  CUDA_TRY (cudaEventRecord (env->rp_ev1, env->stream));
  CUDA_TRY (cudaEventSynchronize (env->rp_ev1));
  float ms = 0.0f;
  CUDA_TRY (cudaEventElapsedTime (&ms, env->rp_ev0, env->rp_ev1));
  env->replay_ms += ms;
  return PRB_OK;
}

/* Environments the parallel engine gave up on (PRB_ERR_TIE: the order of two simultaneous events could not be
 * rebuilt from local information) are replayed by the event-serial engine, whose global (timestamp, uid) order
 * needs no such reconstruction.  Their serial state still sits at the episode's reset, so they are advanced
 * from there to where the parallel launches stand; from then on every read-out first brings them up to date.
 * Called before anything is read from the environments. */
static int
resolve_ties (prb_env *env)
{
  if (env->engine_active != PRB_ENGINE_PARALLEL || !env->d_pblobs) return PRB_OK;
  Static &st = env->dst;
  if (env->log_on || st.policy == PRB_POLICY_NODE_TRACE) return PRB_OK;   /* parity runs: the flag is reported instead */
  unsigned long long n = 0;
  CUDA_TRY (cudaMemcpyAsync (&n, env->d_round_stats + 2, sizeof (n), cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  if (n == 0) return PRB_OK;
  if ((size_t) n != env->tied.size ())
    {
      if (!env->d_tied_flags)
        {
          CUDA_TRY (cudaMalloc (&env->d_tied_flags, (size_t) st.E));
          CUDA_TRY (cudaMalloc (&env->d_tied_list, sizeof (int32_t) * (size_t) st.E));
          env->tied_flag.assign ((size_t) st.E, 0);
        }
      k_tied_flags<<<(st.E + 127) / 128, 128, 0, env->stream>>> (st, env->dps, env->d_pblobs, env->d_tied_flags);
      CUDA_TRY (cudaGetLastError ());
      std::vector<uint8_t> f ((size_t) st.E);
      CUDA_TRY (cudaMemcpyAsync (f.data (), env->d_tied_flags, (size_t) st.E, cudaMemcpyDeviceToHost, env->stream));
      CUDA_TRY (cudaStreamSynchronize (env->stream));
      /* newcomers first catch up with the others (from the reset state to tied_until), then all move together */
      std::vector<int32_t> fresh;
      for (int e = 0; e < st.E; e++)
        if (f[(size_t) e] && !env->tied_flag[(size_t) e])
          {
            fresh.push_back (e);
            env->tied_flag[(size_t) e] = 1;
          }
      if (!fresh.empty () && env->tied_until > 0)
        {
          CUDA_TRY (cudaMemcpyAsync (env->d_tied_list, fresh.data (), sizeof (int32_t) * fresh.size (), cudaMemcpyHostToDevice, env->stream));
          int rc = replay_span (env, (int32_t) fresh.size (), 0, env->tied_until);
          if (rc != PRB_OK) return rc;
        }
      env->tied.insert (env->tied.end (), fresh.begin (), fresh.end ());
      CUDA_TRY (cudaMemcpyAsync (env->d_tied_list, env->tied.data (), sizeof (int32_t) * env->tied.size (), cudaMemcpyHostToDevice, env->stream));
      CUDA_TRY (cudaStreamSynchronize (env->stream));
    }
  if (env->tied_until != env->par_until_arg)
    {
      int rc = replay_span (env, (int32_t) env->tied.size (), env->tied_until, env->par_until_arg);
      if (rc != PRB_OK) return rc;
      env->tied_until = env->par_until_arg;
    }
  return PRB_OK;
}

int
prb_set_engine (prb_env *env, int mode, int team)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  if (mode < PRB_ENGINE_AUTO || mode > PRB_ENGINE_PARALLEL) return fail (PRB_EINVAL, "unknown engine");
  if (team != 0 && team != 8 && team != 16 && team != 32) return fail (PRB_EINVAL, "team size must be 0 (auto), 8, 16 or 32");
  if (mode == PRB_ENGINE_PARALLEL && !env->pm.supported)
    return fail (PRB_EINVAL, "the parallel engine does not support this configuration: " + env->pm.why);
  env->engine_mode = mode;
  if (team) env->team = team;
  return PRB_OK;
}

int prb_engine (prb_env *env) { return env ? env->engine_active : PRB_EINVAL; }

int
prb_round_stats (prb_env *env, int64_t *rounds, int64_t *executed)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  unsigned long long h[2] = { 0, 0 };
  if (env->d_round_stats)
    {
      CUDA_TRY (cudaSetDevice (env->device));
      CUDA_TRY (cudaMemcpyAsync (h, env->d_round_stats, 16, cudaMemcpyDeviceToHost, env->stream));
      CUDA_TRY (cudaStreamSynchronize (env->stream));
    }
  if (rounds) *rounds = (int64_t) h[0];
  if (executed) *executed = (int64_t) h[1];
  return PRB_OK;
}

int
prb_handovers (prb_env *env)
{
  if (!env) return PRB_EINVAL;
  int rc = resolve_ties (env);
  return rc != PRB_OK ? rc : (int) env->tied.size ();
}

int
prb_handover_stats (prb_env *env, int64_t *envs, double *replay_ms)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  int rc = resolve_ties (env);
  if (rc != PRB_OK) return rc;
  if (envs) *envs = (int64_t) env->tied.size ();
  if (replay_ms) *replay_ms = env->replay_ms;
  return PRB_OK;
}

int prb_num_links (prb_env *env) { return env ? env->dst.L : PRB_EINVAL; }

int
prb_state_bytes (prb_env *env, int64_t *blob_bytes, int64_t *pool_bytes)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  bool par = env->engine_active == PRB_ENGINE_PARALLEL;
  if (blob_bytes) *blob_bytes = par ? env->dps.pblob_bytes : env->dst.blob_bytes;
  if (pool_bytes) *pool_bytes = par ? (int64_t) sizeof (PRec) * env->dps.ring_records : (int64_t) sizeof (Pkt) * env->dst.P;
  return PRB_OK;
}

int
prb_set_inject_trace (prb_env *env, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst,
                      const int64_t *uid, const uint8_t *valable, const uint8_t *rejected)
{
  return prb_set_inject_trace_typed (env, n, t_ns, src, dst, uid, valable, rejected, nullptr);
}

int
prb_set_inject_trace_typed (prb_env *env, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst,
                            const int64_t *uid, const uint8_t *valable, const uint8_t *rejected, const uint8_t *ptype)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  CUDA_TRY (cudaSetDevice (env->device));
  Static &st = env->dst;
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  if (n <= 0 || !t_ns)
    {
      st.inject_mode = PRB_INJECT_PHILOX;
      return PRB_OK;
    }
  if (n >= (1ll << 31)) return fail (PRB_EINVAL, "trace too long");
  if (!build_trace (env->hm, n, t_ns, src, dst, uid, valable, rejected, env->tr, ptype))
    return fail (PRB_EINVAL, "injection trace names a flow that the configuration does not have");
  size_t o_t = 0, o_uid = o_t + 8 * n, o_next = o_uid + 4 * n, o_fl = o_next + 4 * n, total = o_fl + n + 16;
  cudaFree (env->d_trace);
  env->d_trace = nullptr;
  CUDA_TRY (cudaMalloc (&env->d_trace, total));
  CUDA_TRY (cudaMemcpyAsync (env->d_trace + o_t, env->tr.t.data (), 8 * n, cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaMemcpyAsync (env->d_trace + o_uid, env->tr.uid.data (), 4 * n, cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaMemcpyAsync (env->d_trace + o_next, env->tr.next.data (), 4 * n, cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaMemcpyAsync (env->d_trace + o_fl, env->tr.flags.data (), n, cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaMemcpyAsync (env->d_flow_first, env->hm.flow_first.data (), sizeof (int32_t) * std::max (st.F, 1),
                             cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  cudaFree (env->d_tr_prev);
  env->d_tr_prev = nullptr;
  CUDA_TRY (cudaMalloc (&env->d_tr_prev, 4 * n + 16));
  CUDA_TRY (cudaMemcpy (env->d_tr_prev, env->tr.prev.data (), 4 * n, cudaMemcpyHostToDevice));
  env->dps.tr_prev = env->d_tr_prev;
  st.inject_mode = PRB_INJECT_TRACE;
  st.tr_n = n;
  st.tr_t = (const int64_t *) (env->d_trace + o_t);
  st.tr_uid = (const uint32_t *) (env->d_trace + o_uid);
  st.tr_next = (const int32_t *) (env->d_trace + o_next);
  st.tr_flags = (const uint8_t *) (env->d_trace + o_fl);
  st.flow_first = env->d_flow_first;
  return PRB_OK;
}

int
prb_set_policy (prb_env *env, int kind, const int32_t *data, int64_t n, uint64_t seed)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  CUDA_TRY (cudaSetDevice (env->device));
  Static &st = env->dst;
  if (kind < PRB_POLICY_ZERO || kind > PRB_POLICY_NODE_TRACE) return fail (PRB_EINVAL, "unknown policy kind");
  if (kind == PRB_POLICY_NODE_TRACE)
    {
      if (!env->pm.supported) return fail (PRB_EINVAL, "per-node action traces need the parallel engine: " + env->pm.why);
      if (!data || n < (int64_t) st.M + 1) return fail (PRB_EINVAL, "a per-node trace is [M+1 offsets][actions]");
      if (data[0] != 0 || (int64_t) data[st.M] != n - (st.M + 1)) return fail (PRB_EINVAL, "per-node trace offsets do not match its length");
    }
  if (kind == PRB_POLICY_TABLE)
    {
      if (!data || n != (int64_t) st.M * st.M) return fail (PRB_EINVAL, "a table policy needs M*M entries");
      /* rows must name existing neighbours or the packet vanishes (data-packet-manager.cc:294-297): allowed */
    }
  if (kind == PRB_POLICY_TRACE && (!data || n <= 0)) return fail (PRB_EINVAL, "a trace policy needs actions");
  if (data && n > 0)
    {
      if (n > env->pol_cap)
        {
          CUDA_TRY (cudaStreamSynchronize (env->stream));
          cudaFree (env->d_pol);
          env->d_pol = nullptr;
          CUDA_TRY (cudaMalloc (&env->d_pol, sizeof (int32_t) * n));
          env->pol_cap = n;
        }
      CUDA_TRY (cudaMemcpyAsync (env->d_pol, data, sizeof (int32_t) * n, cudaMemcpyHostToDevice, env->stream));
    }
  if (kind == PRB_POLICY_TABLE)
    env->pol_host.assign (data, data + n);
  else
    env->pol_host.clear ();
  st.policy = kind;
  st.pol_data = env->d_pol;
  st.pol_n = data ? n : 0;
  env->pol_node_trace = kind == PRB_POLICY_NODE_TRACE;
  if (env->pol_node_trace)
    {
      env->dps.pol_off = env->d_pol;
      st.pol_data = env->d_pol + (st.M + 1);
      st.pol_n = n - (st.M + 1);
    }
  bool reseed = kind == PRB_POLICY_RANDOM || kind == PRB_POLICY_TRACE;
  if (reseed)
    {
      st.policy_seed = seed;
      k_seed_policy<<<(st.E + 127) / 128, 128, 0, env->stream>>> (st, env->d_blobs);
      CUDA_TRY (cudaGetLastError ());
      env->launches++;
    }
  return PRB_OK;
}

int
prb_enable_log (prb_env *env, int64_t event_capacity, int64_t step_capacity)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  CUDA_TRY (cudaSetDevice (env->device));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  Static &st = env->dst;
  cudaFree (env->d_ev); env->d_ev = nullptr;
  cudaFree (env->d_st); env->d_st = nullptr;
  st.ev_cap = std::max<int64_t> (event_capacity, 0);
  st.st_cap = std::max<int64_t> (step_capacity, 0);
  env->log_on = st.ev_cap > 0 || st.st_cap > 0;
  if (st.ev_cap > 0) CUDA_TRY (cudaMalloc (&env->d_ev, sizeof (prb_event) * (size_t) st.ev_cap * st.E));
  if (st.st_cap > 0) CUDA_TRY (cudaMalloc (&env->d_st, sizeof (prb_step_rec) * (size_t) st.st_cap * st.E));
  return PRB_OK;
}

static int
log_counts (prb_env *env, int32_t e, int64_t *ev_n, int64_t *st_n)
{
  const Static &st = env->dst;
  if (e < 0 || e >= st.E) return fail (PRB_EINVAL, "env index out of range");
  if (env->engine_active == PRB_ENGINE_PARALLEL)
    k_log_counts_par<<<1, 1, 0, env->stream>>> (env->dps, env->d_pblobs, e, env->d_scratch);
  else
    k_log_counts<<<1, 1, 0, env->stream>>> (st, env->d_blobs, e, env->d_scratch);
  CUDA_TRY (cudaGetLastError ());
  int64_t h[2];
  CUDA_TRY (cudaMemcpyAsync (h, env->d_scratch, 16, cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  *ev_n = h[0];
  *st_n = h[1];
  return PRB_OK;
}

int
prb_read_events (prb_env *env, int32_t e, prb_event *out, int64_t capacity, int64_t *n)
{
  if (!env || !n) return fail (PRB_EINVAL, "null argument");
  CUDA_TRY (cudaSetDevice (env->device));
  int64_t en, sn;
  int rc = log_counts (env, e, &en, &sn);
  if (rc != PRB_OK) return rc;
  *n = en;
  if (out && en > 0)
    {
      if (!env->d_ev) return fail (PRB_ESTATE, "event log is not enabled");
      int64_t k = std::min (en, capacity);
      CUDA_TRY (cudaMemcpyAsync (out, env->d_ev + (size_t) e * env->dst.ev_cap, sizeof (prb_event) * k, cudaMemcpyDeviceToHost, env->stream));
      CUDA_TRY (cudaStreamSynchronize (env->stream));
    }
  return PRB_OK;
}

int
prb_read_steps (prb_env *env, int32_t e, prb_step_rec *out, int64_t capacity, int64_t *n)
{
  if (!env || !n) return fail (PRB_EINVAL, "null argument");
  CUDA_TRY (cudaSetDevice (env->device));
  int64_t en, sn;
  int rc = log_counts (env, e, &en, &sn);
  if (rc != PRB_OK) return rc;
  *n = sn;
  if (out && sn > 0)
    {
      if (!env->d_st) return fail (PRB_ESTATE, "step log is not enabled");
      int64_t k = std::min (sn, capacity);
      CUDA_TRY (cudaMemcpyAsync (out, env->d_st + (size_t) e * env->dst.st_cap, sizeof (prb_step_rec) * k, cudaMemcpyDeviceToHost, env->stream));
      CUDA_TRY (cudaStreamSynchronize (env->stream));
    }
  return PRB_OK;
}

int
prb_run (prb_env *env, int64_t until_ns)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  CUDA_TRY (cudaSetDevice (env->device));
  Static &st = env->dst;
  /* engine: the node-parallel one whenever the agent is a function of the packet alone and nobody asked for
   * the ordered logs; an episode stays on the engine that started it */
  bool par_policy = st.policy == PRB_POLICY_ZERO || st.policy == PRB_POLICY_TABLE || st.policy == PRB_POLICY_NODE_TRACE;
  /* AUTO: only while an environment's state fits shared memory for the launch — beyond that (hundreds of flows
   * per node: the 100-node fixture holds 159 KB per environment) the event-serial engine is the faster one
   * (measured on B200, 4096 environments: 1.1e8 against 4.1e7 packet-hops/s) */
  /* "fits": at least four environments resident per SM (one per block with ~150 KB of flow timers is one warp per SM) */
  bool par_fits = 4 * ((size_t) env->dps.pblob_bytes + 1024) <= (size_t) env->smem_per_sm;
  bool want_par = env->engine_mode == PRB_ENGINE_PARALLEL
                  || (env->engine_mode == PRB_ENGINE_AUTO && env->pm.supported && par_fits && par_policy && !env->log_on);
  if (env->engine_active == PRB_ENGINE_SERIAL) want_par = false;
  if (env->engine_active == PRB_ENGINE_PARALLEL) want_par = true;
  if (want_par)
    {
      if (!env->pm.supported) return fail (PRB_ESTATE, "the parallel engine does not support this configuration: " + env->pm.why);
      if (!par_policy) return fail (PRB_ESTATE, "the parallel engine runs table / per-node-trace agents only: prb_reset first");
      if (env->d_st) return fail (PRB_ESTATE, "the parallel engine keeps no step log");
      return run_parallel (env, until_ns);
    }
  if (st.policy == PRB_POLICY_NODE_TRACE) return fail (PRB_ESTATE, "per-node action traces need the parallel engine");
  st.mode = MODE_RUN;
  env->engine_active = PRB_ENGINE_SERIAL;
  LaunchArgs a;
  a.blobs = env->d_blobs; a.pools = env->d_pools; a.ev_logs = env->d_ev; a.st_logs = env->d_st;
  a.actions = nullptr; a.out = nullptr; a.until_ns = until_ns;
  a.qw = nullptr; a.steps = 1; a.eps = 0.0f; a.seed = 0; a.step0 = 0; a.replay = ReplayArgs (); a.env_list = nullptr; a.n_list = 0;
  return launch_advance (env, a);
}

int
prb_step (prb_env *env, const int32_t *actions_dev, prb_step_rec *out_dev)
{
  if (!env || !out_dev) return fail (PRB_EINVAL, "null argument");
  CUDA_TRY (cudaSetDevice (env->device));
  Static &st = env->dst;
  if (env->engine_active == PRB_ENGINE_PARALLEL)
    return fail (PRB_ESTATE, "this episode runs on the parallel engine (prb_run with a table agent): prb_reset before prb_step");
  st.mode = MODE_GYM;
  env->engine_active = PRB_ENGINE_SERIAL;
  LaunchArgs a;
  a.blobs = env->d_blobs; a.pools = env->d_pools; a.ev_logs = env->d_ev; a.st_logs = env->d_st;
  a.actions = actions_dev; a.out = out_dev; a.until_ns = st.sim_time_ns;
  a.qw = nullptr; a.steps = 1; a.eps = 0.0f; a.seed = 0; a.step0 = 0; a.replay = ReplayArgs (); a.env_list = nullptr; a.n_list = 0;
  return launch_advance (env, a);
}

int
prb_step_host (prb_env *env, const int32_t *actions, prb_step_rec *out)
{
  if (!env || !out) return fail (PRB_EINVAL, "null argument");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  if (actions)
    CUDA_TRY (cudaMemcpyAsync (env->d_actions, actions, sizeof (int32_t) * (size_t) st.E, cudaMemcpyHostToDevice, env->stream));
  int rc = prb_step (env, env->d_actions, env->d_out);
  if (rc != PRB_OK) return rc;
  CUDA_TRY (cudaMemcpyAsync (out, env->d_out, sizeof (prb_step_rec) * (size_t) st.E, cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}


int64_t
prb_qnet_num_floats (prb_env *env)
{
  if (!env) return PRB_EINVAL;
  return qnet_floats_per_node (env->dst.M, env->dst.K) * env->dst.M;
}

int
prb_qnet_load (prb_env *env, const float *weights, int64_t n_floats)
{
  if (!env || !weights) return fail (PRB_EINVAL, "null argument");
  CUDA_TRY (cudaSetDevice (env->device));
  if (n_floats != prb_qnet_num_floats (env)) return fail (PRB_EINVAL, "weight count differs from prb_qnet_num_floats");
  if (!env->d_qw) CUDA_TRY (cudaMalloc (&env->d_qw, sizeof (float) * (size_t) n_floats));
  CUDA_TRY (cudaMemcpyAsync (env->d_qw, weights, sizeof (float) * (size_t) n_floats, cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}

/* the pending decisions of all environments, bucketed by node and evaluated on the tensor cores (k_qnet_mma);
 * ra.rows != null: the replay bookkeeping of every record happens here too */
static int
qnet_act_batched (prb_env *env, const prb_step_rec *rec_dev, int32_t *actions_dev, float *q_out_dev, float eps, uint64_t seed,
                  const ReplayArgs &ra)
{
  const Static &st = env->dst;
  if (!env->d_qcount)
    {
      CUDA_TRY (cudaMalloc (&env->d_qcount, sizeof (int32_t) * 2 * (size_t) (st.M + 1)));
      CUDA_TRY (cudaMalloc (&env->d_qperm, sizeof (int32_t) * (size_t) st.M * (size_t) st.E));
      CUDA_TRY (cudaMemsetAsync (env->d_qcount, 0, sizeof (int32_t) * 2 * (size_t) (st.M + 1), env->stream));
      env->q_parity = 0;
    }
  QBatchArgs qb;
  qb.count = env->d_qcount; qb.perm = env->d_qperm; qb.parity = env->q_parity;
  env->q_parity ^= 1;
  k_qnet_bucket<<<(st.E + 127) / 128, 128, 0, env->stream>>> (st, ra, rec_dev, actions_dev, qb);
  CUDA_TRY (cudaGetLastError ());
  const int tiles = st.E / 16 + st.M + 1;
  k_qnet_mma<<<(tiles + QM_WARPS - 1) / QM_WARPS, QM_WARPS * 32, 0, env->stream>>> (st, ra, env->d_qw, rec_dev, actions_dev, q_out_dev, eps, seed,
                                                                                    env->q_step++, env->rp.target_mode ? env->rp.tmin : nullptr, qb);
  CUDA_TRY (cudaGetLastError ());
  env->launches += 2;
  return PRB_OK;
}

int
prb_qnet_set_arch (prb_env *env, int hidden_layers)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  if (hidden_layers != 1 && hidden_layers != 2) return fail (PRB_EINVAL, "the Q-networks have 1 or 2 hidden layers after the concatenation");
  env->dst.q_hidden = hidden_layers;
  return PRB_OK;
}

int
prb_qnet_set_mode (prb_env *env, int mode)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  if (mode < PRB_QNET_FUSED || mode > PRB_QNET_MMA) return fail (PRB_EINVAL, "unknown Q-network mode");
  env->qnet_mode = mode;
  return PRB_OK;
}

int
prb_qnet_act (prb_env *env, const prb_step_rec *rec_dev, int32_t *actions_dev, float *q_out_dev, float eps, uint64_t seed)
{
  if (!env || !rec_dev || !actions_dev) return fail (PRB_EINVAL, "null argument");
  if (!env->d_qw) return fail (PRB_ESTATE, "no Q-network loaded (prb_qnet_load)");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  if (env->qnet_mode == PRB_QNET_MMA) return qnet_act_batched (env, rec_dev, actions_dev, q_out_dev, eps, seed, ReplayArgs ());
  k_qnet_act<<<(st.E + 3) / 4, 128, 0, env->stream>>> (st, env->d_qw, rec_dev, actions_dev, q_out_dev, eps, seed, env->q_step++,
                                                       env->rp.target_mode ? env->rp.tmin : nullptr);
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  return PRB_OK;
}

int
prb_replay_enable (prb_env *env, int64_t rows_per_node, int32_t pending_slots)
{
  if (!env || rows_per_node < 1 || pending_slots < 1 || (pending_slots & (pending_slots - 1))) return fail (PRB_EINVAL, "bad argument");
  if (env->rp.rows) return fail (PRB_ESTATE, "replay buffers already enabled");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  ReplayArgs &ra = env->rp;
  ra.cap = rows_per_node; ra.pcap = pending_slots;
  ra.R = 2 * (1 + st.K) + 3;
  ra.PW = 4 + 1 + st.K;
  size_t rows_b = sizeof (float) * (size_t) st.M * (size_t) ra.cap * ra.R;
  size_t pend_b = sizeof (int32_t) * (size_t) st.E * (size_t) ra.pcap * ra.PW;
  CUDA_TRY (cudaMalloc (&ra.rows, rows_b));
  CUDA_TRY (cudaMalloc (&ra.cursor, sizeof (unsigned long long) * (st.M + 1)));
  CUDA_TRY (cudaMalloc (&ra.pending, pend_b));
  ra.missed = ra.cursor + st.M;
  CUDA_TRY (cudaMalloc (&ra.tmin, sizeof (float) * (size_t) st.E));
  CUDA_TRY (cudaMemsetAsync (ra.tmin, 0x7F, sizeof (float) * (size_t) st.E, env->stream));    /* 3.39e38: "none" until written */
  CUDA_TRY (cudaMemsetAsync (ra.rows, 0, rows_b, env->stream));
  CUDA_TRY (cudaMemsetAsync (ra.cursor, 0, sizeof (unsigned long long) * (st.M + 1), env->stream));
  CUDA_TRY (cudaMemsetAsync (ra.pending, 0xFF, pend_b, env->stream));
  return PRB_OK;
}

int
prb_replay_set_target_mode (prb_env *env, int on, float gamma)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  if (!env->rp.rows) return fail (PRB_ESTATE, "replay buffers not enabled (prb_replay_enable)");
  env->rp.target_mode = on ? 1 : 0;
  env->rp.gamma = gamma;
  return PRB_OK;
}

int
prb_replay_record (prb_env *env, const prb_step_rec *rec_dev, const int32_t *actions_dev)
{
  if (!env || !rec_dev || !actions_dev) return fail (PRB_EINVAL, "null argument");
  if (!env->rp.rows) return fail (PRB_ESTATE, "replay buffers not enabled (prb_replay_enable)");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  k_replay_record<<<(st.E + 127) / 128, 128, 0, env->stream>>> (st, env->rp, rec_dev, actions_dev);
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  return PRB_OK;
}

int
prb_replay_counts (prb_env *env, int64_t *rows_written /* host [M] */, int64_t *missed /* host */)
{
  if (!env || !rows_written) return fail (PRB_EINVAL, "null argument");
  if (!env->rp.rows) return fail (PRB_ESTATE, "replay buffers not enabled (prb_replay_enable)");
  CUDA_TRY (cudaSetDevice (env->device));
  std::vector<unsigned long long> h (env->dst.M + 1);
  CUDA_TRY (cudaMemcpyAsync (h.data (), env->rp.cursor, sizeof (unsigned long long) * h.size (), cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  for (int m = 0; m < env->dst.M; m++) rows_written[m] = (int64_t) h[m];
  if (missed) *missed = (int64_t) h[env->dst.M];
  return PRB_OK;
}

int
prb_replay_read (prb_env *env, int32_t node, float *rows_host, int64_t max_rows, int64_t *n_rows)
{
  if (!env || !n_rows) return fail (PRB_EINVAL, "null argument");
  if (!env->rp.rows) return fail (PRB_ESTATE, "replay buffers not enabled (prb_replay_enable)");
  if (node < 0 || node >= env->dst.M) return fail (PRB_EINVAL, "node out of range");
  CUDA_TRY (cudaSetDevice (env->device));
  unsigned long long w = 0;
  CUDA_TRY (cudaMemcpyAsync (&w, env->rp.cursor + node, sizeof (w), cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  int64_t n = (int64_t) std::min<unsigned long long> (w, (unsigned long long) env->rp.cap);
  *n_rows = n;
  if (rows_host && max_rows > 0)
    {
      n = std::min<int64_t> (n, max_rows);
      CUDA_TRY (cudaMemcpyAsync (rows_host, env->rp.rows + (size_t) node * env->rp.cap * env->rp.R, sizeof (float) * (size_t) n * env->rp.R,
                                 cudaMemcpyDeviceToHost, env->stream));
      CUDA_TRY (cudaStreamSynchronize (env->stream));
    }
  return PRB_OK;
}

int
prb_replay_sample (prb_env *env, int32_t node, int32_t batch, uint64_t seed, float *obs_dev, int32_t *action_dev, float *reward_dev,
                   float *next_obs_dev, float *done_dev)
{
  if (!env || !obs_dev || !action_dev || !reward_dev || !next_obs_dev || !done_dev || batch < 1) return fail (PRB_EINVAL, "bad argument");
  if (!env->rp.rows) return fail (PRB_ESTATE, "replay buffers not enabled (prb_replay_enable)");
  if (node < 0 || node >= env->dst.M) return fail (PRB_EINVAL, "node out of range");
  CUDA_TRY (cudaSetDevice (env->device));
  k_replay_sample<<<(batch + 127) / 128, 128, 0, env->stream>>> (env->rp, node, batch, 1 + env->dst.K, seed, env->rp_call++, obs_dev, action_dev,
                                                                  reward_dev, next_obs_dev, done_dev);
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  return PRB_OK;
}

int
prb_replay_enable_prioritized (prb_env *env, float alpha)
{
  if (!env || alpha < 0.0f) return fail (PRB_EINVAL, "bad argument");
  if (!env->rp.rows) return fail (PRB_ESTATE, "replay buffers not enabled (prb_replay_enable)");
  if (env->rp.row_gstep) return fail (PRB_ESTATE, "prioritized replay already enabled");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  ReplayArgs &ra = env->rp;
  size_t rows = (size_t) st.M * (size_t) ra.cap;
  std::vector<int32_t> ones ((size_t) st.M * st.K + st.M, 1);       /* latest_gradient_step = [1] * nb_neighbors (:411) */
  int32_t *ints = nullptr;
  CUDA_TRY (cudaMalloc (&ra.row_gstep, sizeof (float) * rows));
  CUDA_TRY (cudaMalloc (&ints, sizeof (int32_t) * ones.size ()));
  CUDA_TRY (cudaMalloc (&ra.psum, sizeof (double)));
  CUDA_TRY (cudaMemsetAsync (ra.row_gstep, 0, sizeof (float) * rows, env->stream));
  CUDA_TRY (cudaMemcpyAsync (ints, ones.data (), sizeof (int32_t) * ones.size (), cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  ra.latest = ints;
  ra.node_gstep = ints + (size_t) st.M * st.K;
  ra.alpha = alpha;
  return PRB_OK;
}

int
prb_replay_set_gradient_steps (prb_env *env, const int32_t *steps /* host [M] */)
{
  if (!env || !steps) return fail (PRB_EINVAL, "null argument");
  if (!env->rp.row_gstep) return fail (PRB_ESTATE, "prioritized replay not enabled (prb_replay_enable_prioritized)");
  CUDA_TRY (cudaSetDevice (env->device));
  CUDA_TRY (cudaMemcpyAsync (const_cast<int32_t *> (env->rp.node_gstep), steps, sizeof (int32_t) * env->dst.M, cudaMemcpyHostToDevice,
                             env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}

int
prb_replay_sample_weighted (prb_env *env, int32_t node, int32_t batch, uint64_t seed, float *obs_dev, int32_t *action_dev,
                            float *reward_dev, float *next_obs_dev, float *done_dev, float *weights_dev)
{
  if (!env || !obs_dev || !action_dev || !reward_dev || !next_obs_dev || !done_dev || !weights_dev || batch < 1)
    return fail (PRB_EINVAL, "bad argument");
  if (!env->rp.rows) return fail (PRB_ESTATE, "replay buffers not enabled (prb_replay_enable)");
  if (node < 0 || node >= env->dst.M) return fail (PRB_EINVAL, "node out of range");
  CUDA_TRY (cudaSetDevice (env->device));
  const int ow = 1 + env->dst.K;
  if (env->rp.row_gstep)
    {
      CUDA_TRY (cudaMemsetAsync (env->rp.psum, 0, sizeof (double), env->stream));
      k_replay_priority_sum<<<64, 256, 0, env->stream>>> (env->rp, node, ow, env->dst.K);
      env->launches++;
    }
  k_replay_sample<<<(batch + 127) / 128, 128, 0, env->stream>>> (env->rp, node, batch, ow, seed, env->rp_call++, obs_dev, action_dev,
                                                                  reward_dev, next_obs_dev, done_dev, weights_dev, env->dst.K);
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  return PRB_OK;
}

/* ---- learner on the device ---- */
int
prb_replay_sample_all (prb_env *env, int32_t batch, uint64_t seed, int64_t min_rows, float *obs_dev, int32_t *action_dev, float *reward_dev,
                       float *next_obs_dev, float *done_dev, float *weights_dev, float *active_dev)
{
  if (!env || !obs_dev || !action_dev || !reward_dev || !next_obs_dev || !done_dev || !active_dev || batch < 1)
    return fail (PRB_EINVAL, "bad argument");
  if (!env->rp.rows) return fail (PRB_ESTATE, "replay buffers not enabled (prb_replay_enable)");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  const int ow = 1 + st.K;
  std::vector<unsigned long long> cur ((size_t) st.M);
  CUDA_TRY (cudaMemcpyAsync (cur.data (), env->rp.cursor, sizeof (unsigned long long) * st.M, cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  std::vector<float> act ((size_t) st.M, 0.0f);
  for (int m = 0; m < st.M; m++)
    {
      const int64_t have = (int64_t) std::min<unsigned long long> (cur[(size_t) m], (unsigned long long) env->rp.cap);
      if (have < std::max<int64_t> (min_rows, 1)) continue;
      act[(size_t) m] = 1.0f;
      const size_t o = (size_t) m * batch;
      float *wts = weights_dev ? weights_dev + o : nullptr;
      if (env->rp.row_gstep && wts)
        {
          CUDA_TRY (cudaMemsetAsync (env->rp.psum, 0, sizeof (double), env->stream));
          k_replay_priority_sum<<<64, 256, 0, env->stream>>> (env->rp, m, ow, st.K);
          env->launches++;
        }
      k_replay_sample<<<(batch + 127) / 128, 128, 0, env->stream>>> (env->rp, m, batch, ow, seed, env->rp_call, obs_dev + o * ow, action_dev + o,
                                                                      reward_dev + o, next_obs_dev + o * ow, done_dev + o, wts, st.K);
      CUDA_TRY (cudaGetLastError ());
      env->launches++;
    }
  env->rp_call++;
  CUDA_TRY (cudaMemcpyAsync (active_dev, act.data (), sizeof (float) * st.M, cudaMemcpyHostToDevice, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));        /* `act` leaves scope */
  return PRB_OK;
}

int
prb_learner_init (prb_env *env, float lr, float gamma, float beta1, float beta2, float adam_eps, int target_mode)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  if (!env->d_qw) return fail (PRB_ESTATE, "no Q-network loaded (prb_qnet_load): the learner trains the loaded weights in place");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  const size_t n = (size_t) st.M * (size_t) qnet_floats_per_node (st.M, st.K);
  if (!env->d_lw_target)
    {
      CUDA_TRY (cudaMalloc (&env->d_lw_target, sizeof (float) * n));
      CUDA_TRY (cudaMalloc (&env->d_lexp, sizeof (float) * n));
      CUDA_TRY (cudaMalloc (&env->d_lexp_sq, sizeof (float) * n));
      CUDA_TRY (cudaMalloc (&env->d_lsteps, sizeof (float) * st.M));
      CUDA_TRY (cudaMalloc (&env->d_ltd, sizeof (float) * st.M));
    }
  CUDA_TRY (cudaMemcpyAsync (env->d_lw_target, env->d_qw, sizeof (float) * n, cudaMemcpyDeviceToDevice, env->stream));
  CUDA_TRY (cudaMemsetAsync (env->d_lexp, 0, sizeof (float) * n, env->stream));
  CUDA_TRY (cudaMemsetAsync (env->d_lexp_sq, 0, sizeof (float) * n, env->stream));
  CUDA_TRY (cudaMemsetAsync (env->d_lsteps, 0, sizeof (float) * st.M, env->stream));
  env->l_lr = lr; env->l_gamma = gamma; env->l_b1 = beta1; env->l_b2 = beta2; env->l_eps = adam_eps; env->l_target_mode = target_mode ? 1 : 0;
  return PRB_OK;
}

int
prb_learner_sync_targets (prb_env *env)
{
  if (!env || !env->d_lw_target) return fail (PRB_ESTATE, "no learner (prb_learner_init)");
  CUDA_TRY (cudaSetDevice (env->device));
  const size_t n = (size_t) env->dst.M * (size_t) qnet_floats_per_node (env->dst.M, env->dst.K);
  CUDA_TRY (cudaMemcpyAsync (env->d_lw_target, env->d_qw, sizeof (float) * n, cudaMemcpyDeviceToDevice, env->stream));
  return PRB_OK;
}

int
prb_learner_grad (prb_env *env, int32_t batch, const float *obs_dev, const int32_t *action_dev, const float *reward_dev,
                  const float *next_obs_dev, const float *done_dev, const float *weights_dev, const float *active_dev, float *bucket_dev,
                  float *td_mse_host)
{
  if (!env || !obs_dev || !action_dev || !reward_dev || !next_obs_dev || !done_dev || !active_dev || !bucket_dev || batch < 1)
    return fail (PRB_EINVAL, "bad argument");
  if (!env->d_lw_target) return fail (PRB_ESTATE, "no learner (prb_learner_init)");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  const int P = (int) qnet_floats_per_node (st.M, st.K);
  const size_t smem = sizeof (float) * (size_t) P;
  if (smem > (size_t) env->max_smem) return fail (PRB_EINVAL, "one node's parameters do not fit shared memory");
  CUDA_TRY (cudaMemsetAsync (bucket_dev, 0, sizeof (float) * (size_t) st.M * P, env->stream));
  CUDA_TRY (cudaMemcpyAsync (bucket_dev + (size_t) st.M * P, active_dev, sizeof (float) * st.M, cudaMemcpyDeviceToDevice, env->stream));
  CUDA_TRY (cudaMemsetAsync (env->d_ltd, 0, sizeof (float) * st.M, env->stream));
  LearnArgs a;
  a.W = env->d_qw; a.Wt = env->d_lw_target; a.obs = obs_dev; a.nxt = next_obs_dev; a.act = action_dev; a.rew = reward_dev; a.done = done_dev;
  a.wts = weights_dev; a.active = active_dev; a.grad = bucket_dev; a.td_sq = env->d_ltd; a.B = batch; a.target_mode = env->l_target_mode;
  a.gamma = env->l_gamma;
  CUDA_TRY (cudaFuncSetAttribute (k_learn_grad, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
  const int chunks = (batch + LEARN_ROWS - 1) / LEARN_ROWS;
  k_learn_grad<<<st.M * chunks, 32, smem, env->stream>>> (st, a);
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  if (td_mse_host)
    {
      CUDA_TRY (cudaMemcpyAsync (td_mse_host, env->d_ltd, sizeof (float) * st.M, cudaMemcpyDeviceToHost, env->stream));
      CUDA_TRY (cudaStreamSynchronize (env->stream));
      for (int m = 0; m < st.M; m++) td_mse_host[m] /= (float) batch;
    }
  return PRB_OK;
}

int
prb_learner_apply (prb_env *env, const float *bucket_dev, int32_t world_size)
{
  if (!env || !bucket_dev || world_size < 1) return fail (PRB_EINVAL, "bad argument");
  if (!env->d_lw_target) return fail (PRB_ESTATE, "no learner (prb_learner_init)");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  const int P = (int) qnet_floats_per_node (st.M, st.K);
  dim3 grid ((P + 255) / 256, st.M);
  k_learn_adam<<<grid, 256, 0, env->stream>>> (env->d_qw, env->d_lexp, env->d_lexp_sq, env->d_lsteps, bucket_dev, st.M, P, 1.0f / (float) world_size,
                                               env->l_lr, env->l_b1, env->l_b2, env->l_eps);
  CUDA_TRY (cudaGetLastError ());
  k_learn_bump<<<(st.M + 31) / 32, 32, 0, env->stream>>> (env->d_lsteps, bucket_dev, st.M, P);
  CUDA_TRY (cudaGetLastError ());
  env->launches += 2;
  return PRB_OK;
}

int
prb_learner_read (prb_env *env, int32_t which, float *out_host, int64_t n_floats)
{
  if (!env || !out_host) return fail (PRB_EINVAL, "null argument");
  if (!env->d_lw_target) return fail (PRB_ESTATE, "no learner (prb_learner_init)");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  const int64_t n = (int64_t) st.M * qnet_floats_per_node (st.M, st.K);
  const float *src = which == 0 ? env->d_qw : (which == 1 ? env->d_lw_target : (which == 2 ? env->d_lexp : (which == 3 ? env->d_lexp_sq : env->d_lsteps)));
  const int64_t have = which == 4 ? st.M : n;
  if (which < 0 || which > 4 || n_floats != have) return fail (PRB_EINVAL, "which is 0..4 (weights, target, exp_avg, exp_avg_sq, steps) with its float count");
  CUDA_TRY (cudaMemcpyAsync (out_host, src, sizeof (float) * (size_t) have, cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}

int
prb_replay_read_priorities (prb_env *env, int32_t node, float *gstep_host, int64_t max_rows, int32_t *latest_host /* [K] */)
{
  if (!env || !gstep_host || !latest_host) return fail (PRB_EINVAL, "null argument");
  if (!env->rp.row_gstep) return fail (PRB_ESTATE, "prioritized replay not enabled (prb_replay_enable_prioritized)");
  if (node < 0 || node >= env->dst.M) return fail (PRB_EINVAL, "node out of range");
  CUDA_TRY (cudaSetDevice (env->device));
  int64_t n = std::min<int64_t> (max_rows, env->rp.cap);
  CUDA_TRY (cudaMemcpyAsync (gstep_host, env->rp.row_gstep + (size_t) node * env->rp.cap, sizeof (float) * (size_t) n, cudaMemcpyDeviceToHost,
                             env->stream));
  CUDA_TRY (cudaMemcpyAsync (latest_host, env->rp.latest + (size_t) node * env->dst.K, sizeof (int32_t) * env->dst.K,
                             cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}

int
prb_step_qnet (prb_env *env, int32_t n_steps, float eps, uint64_t seed, prb_step_rec *out_dev)
{
  if (!env || !out_dev || n_steps < 1) return fail (PRB_EINVAL, "bad argument");
  if (!env->d_qw) return fail (PRB_ESTATE, "no Q-network loaded (prb_qnet_load)");
  int rc = PRB_OK;
  if (env->qnet_mode == PRB_QNET_MMA)
    {
      /* two launches + the simulator's per step: decisions of all environments bucketed by node, Q-networks on the
       * tensor cores, replay bookkeeping in the same kernel */
      CUDA_TRY (cudaSetDevice (env->device));
      for (int i = 0; i < n_steps && rc == PRB_OK; i++)
        {
          rc = qnet_act_batched (env, out_dev, env->d_actions, nullptr, eps, seed, env->rp);
          if (rc == PRB_OK) rc = prb_step (env, env->d_actions, out_dev);
        }
      return rc;
    }
  /* the fused kernel holds the two-hidden-layer network only: one-hidden-layer models take the launch sequence below */
  if (env->G == 32 && !env->qnet_unfused && env->qnet_mode == PRB_QNET_FUSED && env->dst.q_hidden == 2)
    {
      /* one launch: every warp answers the steps of its environment itself (k_advance, gym mode) */
      CUDA_TRY (cudaSetDevice (env->device));
      Static &st = env->dst;
      if (env->engine_active == PRB_ENGINE_PARALLEL)
        return fail (PRB_ESTATE, "this episode runs on the parallel engine (prb_run with a table agent): prb_reset before prb_step");
      st.mode = MODE_GYM;
      env->engine_active = PRB_ENGINE_SERIAL;
      LaunchArgs a;
      a.blobs = env->d_blobs; a.pools = env->d_pools; a.ev_logs = env->d_ev; a.st_logs = env->d_st;
      a.actions = nullptr; a.out = out_dev; a.until_ns = st.sim_time_ns;
      a.qw = env->d_qw; a.steps = n_steps; a.eps = eps; a.seed = seed; a.step0 = env->q_step; a.replay = env->rp;
      a.env_list = nullptr; a.n_list = 0;
      env->q_step += (uint64_t) n_steps;
      return launch_advance (env, a);
    }
  for (int i = 0; i < n_steps && rc == PRB_OK; i++)
    {
      /* the first call of an episode has no pending step: every record is invalid and the actions are zeros */
      rc = prb_qnet_act (env, out_dev, env->d_actions, nullptr, eps, seed);
      if (rc == PRB_OK && env->rp.rows) rc = prb_replay_record (env, out_dev, env->d_actions);
      if (rc == PRB_OK) rc = prb_step (env, env->d_actions, out_dev);
    }
  return rc;
}

int
prb_counters_dev (prb_env *env, prb_counters *out_dev)
{
  if (!env || !out_dev) return fail (PRB_EINVAL, "null argument");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  int rc = resolve_ties (env);
  if (rc != PRB_OK) return rc;
  if (env->engine_active == PRB_ENGINE_PARALLEL)
    {
      k_counters_par<<<(st.E + 127) / 128, 128, 0, env->stream>>> (st, env->dps, env->d_pblobs, out_dev);
      if (!env->tied.empty ())
        k_counters_list<<<((int) env->tied.size () + 127) / 128, 128, 0, env->stream>>> (st, env->d_blobs, env->d_tied_list, (int) env->tied.size (),
                                                                                       out_dev);
    }
  else
    k_counters<<<(st.E + 127) / 128, 128, 0, env->stream>>> (st, env->d_blobs, out_dev);
  CUDA_TRY (cudaGetLastError ());
  env->launches++;
  return PRB_OK;
}

int
prb_counters_host (prb_env *env, prb_counters *out)
{
  if (!env || !out) return fail (PRB_EINVAL, "null argument");
  int rc = prb_counters_dev (env, env->d_counters);
  if (rc != PRB_OK) return rc;
  CUDA_TRY (cudaMemcpyAsync (out, env->d_counters, sizeof (prb_counters) * (size_t) env->dst.E, cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}

int
prb_queue_bytes (prb_env *env, int32_t e, int32_t *out)
{
  if (!env || !out) return fail (PRB_EINVAL, "null argument");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  if (e < 0 || e >= st.E) return fail (PRB_EINVAL, "env index out of range");
  if (st.L > 1000) return fail (PRB_EINVAL, "too many links for the scratch buffer");
  int rc = resolve_ties (env);
  if (rc != PRB_OK) return rc;
  if (env->engine_active == PRB_ENGINE_PARALLEL && !(e < (int) env->tied_flag.size () && env->tied_flag[(size_t) e]))
    k_queue_bytes_par<<<(st.L + 127) / 128, 128, 0, env->stream>>> (st, env->dps, env->d_pblobs, e, (int32_t *) env->d_scratch);
  else
    k_queue_bytes<<<(st.L + 127) / 128, 128, 0, env->stream>>> (st, env->d_blobs, e, (int32_t *) env->d_scratch);
  CUDA_TRY (cudaGetLastError ());
  CUDA_TRY (cudaMemcpyAsync (out, env->d_scratch, sizeof (int32_t) * st.L, cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}

int
prb_queue_bytes_noalloc (prb_env *env, int32_t e, int32_t *out, int32_t out_capacity)
{
  if (!env || !out) return fail (PRB_EINVAL, "null argument");
  CUDA_TRY (cudaSetDevice (env->device));
  const Static &st = env->dst;
  if (e < 0 || e >= st.E) return fail (PRB_EINVAL, "env index out of range");
  if (st.L > PRB_NOALLOC_MAX_LINKS)
    return fail (PRB_EINVAL, "prb_queue_bytes_noalloc: " + std::to_string (st.L) + " links exceed the static scratch of "
                                 + std::to_string (PRB_NOALLOC_MAX_LINKS) + " (use prb_queue_bytes)");
  if (out_capacity < st.L) return fail (PRB_EINVAL, "prb_queue_bytes_noalloc: the output holds fewer than L entries");
  int32_t *scratch = env->device < PRB_NOALLOC_MAX_DEVICES ? g_noalloc_ptr[env->device] : nullptr;
  if (!scratch) return fail (PRB_ESTATE, "prb_queue_bytes_noalloc: no static scratch on this device (device index beyond the table)");
  bool tied_env = false;
  if (env->engine_active == PRB_ENGINE_PARALLEL && env->d_pblobs && !(env->log_on || st.policy == PRB_POLICY_NODE_TRACE))
    {
      /* hand-overs are brought up to date by calls that may allocate (prb_handovers): this one only checks */
      unsigned long long n = 0;
      CUDA_TRY (cudaMemcpyAsync (&n, env->d_round_stats + 2, sizeof (n), cudaMemcpyDeviceToHost, env->stream));
      CUDA_TRY (cudaStreamSynchronize (env->stream));
      if ((size_t) n != env->tied.size () || (n > 0 && env->tied_until != env->par_until_arg))
        return fail (PRB_ESTATE, "prb_queue_bytes_noalloc: environments handed over to the serial engine are pending; call prb_handovers first");
      tied_env = e < (int) env->tied_flag.size () && env->tied_flag[(size_t) e];
    }
  CUDA_TRY (cudaMemsetAsync (scratch, 0, sizeof (int32_t) * (size_t) st.L, env->stream));   /* the array keeps the last call's contents */
  if (env->engine_active == PRB_ENGINE_PARALLEL && !tied_env)
    k_queue_bytes_par<<<(st.L + 127) / 128, 128, 0, env->stream>>> (st, env->dps, env->d_pblobs, e, scratch);
  else
    k_queue_bytes<<<(st.L + 127) / 128, 128, 0, env->stream>>> (st, env->d_blobs, e, scratch);
  CUDA_TRY (cudaGetLastError ());
  CUDA_TRY (cudaMemcpyAsync (out, scratch, sizeof (int32_t) * st.L, cudaMemcpyDeviceToHost, env->stream));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}

int
prb_sync (prb_env *env)
{
  if (!env) return fail (PRB_EINVAL, "null handle");
  CUDA_TRY (cudaSetDevice (env->device));
  CUDA_TRY (cudaStreamSynchronize (env->stream));
  return PRB_OK;
}

int64_t prb_launch_count (prb_env *env) { return env ? env->launches : 0; }

}  /* extern "C" */
