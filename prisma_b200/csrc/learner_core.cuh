/* prisma_b200 — the DQN learner's gradient step on the device (learner half).
 *
 * Mirrors DQN_AGENT.train / get_neighbor_target_value (prisma/source/learner.py:162-189,231-255) and Trainer.step
 * (prisma/source/trainer.py:45-99) for the dqn_buffer agents: node i holds a batch (s, a, r, s', done[, weight]); the TD
 * target of a transition whose action pointed at neighbour j is r + gamma * (1 - done) * min over the actions of j that
 * do not lead back to i of Q_target_j (s') ("target" signalling: the stored reward already is the target); the loss is the
 * mean Huber error of Q_i (s)[a] against it; the optimiser is Adam with Keras' defaults, one state per node.
 * k_learn_grad: one warp per block walks 16 transitions of one node: target, forward with the activations kept in
 * registers, backward by hand (elu' = out > 0 ? 1 : out + 1), gradients summed in a shared-memory image of the node's
 * parameter vector and added to the global bucket at the end.  k_learn_adam applies the step after the (optional)
 * all-reduce of the bucket.  Included by prisma_b200.cu after qnet_forward's helpers. */
#pragma once

struct LearnArgs
{
  const float *W, *Wt;            /* [M][P] online / target parameters (layout of prb_qnet_load) */
  const float *obs, *nxt;         /* [M][B][1 + K] */
  const int32_t *act;             /* [M][B] */
  const float *rew, *done, *wts;  /* [M][B]; wts may be null (uniform replay) */
  const float *active;            /* [M]: 1 when the node has a batch */
  float *grad;                    /* [M][P], zeroed by the caller */
  float *td_sq;                   /* [M] sum of squared TD errors, zeroed by the caller */
  int B, target_mode;
  float gamma;
};

struct QPtr
{
  const float *w_dst, *b_dst, *w_buf, *b_buf, *w1, *b1, *w2, *b2, *w3, *b3;
};

__device__ __forceinline__ QPtr
q_ptrs (const float *W, int M, int K)
{
  QPtr p;
  p.w_dst = W; p.b_dst = p.w_dst + (size_t) M * QH1; p.w_buf = p.b_dst + QH1; p.b_buf = p.w_buf + (size_t) K * QH1;
  p.w1 = p.b_buf + QH1; p.b1 = p.w1 + 2 * QH1 * QH2; p.w2 = p.b1 + QH2; p.b2 = p.w2 + QH2 * QH2; p.w3 = p.b2 + QH2;
  p.b3 = p.w3 + (size_t) QH2 * K;
  return p;
}

struct QRow
{
  float hd, hb, h1a, h1b, h2a, h2b, q, xn;     /* per lane: unit `lane` (and lane + 32) of every layer */
  int dst;
};

/* forward of one observation by a warp, same operation order as qnet_forward; o = [dst, q_1 .. q_k, 0 ...] as floats */
__device__ __forceinline__ void
q_forward_row (const QPtr &p, int M, int K, int k, const float *o, int lane, QRow &r, int hidden)
{
  int dst = (int) o[0];
  dst = dst < 0 ? 0 : (dst >= M ? M - 1 : dst);
  r.dst = dst;
  r.hd = elu_ (p.w_dst[(size_t) dst * QH1 + lane] + p.b_dst[lane]);
  float x = lane < k ? o[1 + lane] : 0.0f;
  float sum = x;
  for (int s = 16; s > 0; s >>= 1) sum += __shfl_xor_sync (0xFFFFFFFFu, sum, s);
  const float mean = sum / (float) k;
  const float dv = lane < k ? x - mean : 0.0f;
  float var = dv * dv;
  for (int s = 16; s > 0; s >>= 1) var += __shfl_xor_sync (0xFFFFFFFFu, var, s);
  r.xn = dv * rsqrtf (var / (float) k + 1e-3f);
  float acc = p.b_buf[lane];
  for (int i = 0; i < k; i++) acc = fmaf (__shfl_sync (0xFFFFFFFFu, r.xn, i), p.w_buf[(size_t) i * QH1 + lane], acc);
  r.hb = elu_ (acc);
  float a0 = p.b1[lane], a1 = p.b1[lane + 32];
#pragma unroll 8
  for (int i = 0; i < 64; i++)
    {
      const float c = __shfl_sync (0xFFFFFFFFu, i < 32 ? r.hd : r.hb, i & 31);
      a0 = fmaf (c, p.w1[(size_t) i * QH2 + lane], a0);
      a1 = fmaf (c, p.w1[(size_t) i * QH2 + lane + 32], a1);
    }
  r.h1a = elu_ (a0); r.h1b = elu_ (a1);
  if (hidden > 1)
    {
      a0 = p.b2[lane]; a1 = p.b2[lane + 32];
#pragma unroll 8
      for (int i = 0; i < 64; i++)
        {
          const float c = __shfl_sync (0xFFFFFFFFu, i < 32 ? r.h1a : r.h1b, i & 31);
          a0 = fmaf (c, p.w2[(size_t) i * QH2 + lane], a0);
          a1 = fmaf (c, p.w2[(size_t) i * QH2 + lane + 32], a1);
        }
      r.h2a = elu_ (a0); r.h2b = elu_ (a1);
    }
  else
    {
      r.h2a = r.h1a; r.h2b = r.h1b;        /* one hidden layer: the output layer reads it directly */
    }
  float q = lane < k ? p.b3[lane] : 0.0f;
#pragma unroll 8
  for (int i = 0; i < 64; i++)
    {
      const float c = __shfl_sync (0xFFFFFFFFu, i < 32 ? r.h2a : r.h2b, i & 31);
      if (lane < k) q = fmaf (c, p.w3[(size_t) i * K + lane], q);
    }
  r.q = lane < k ? elu_ (q) : 0.0f;
}

__device__ __forceinline__ float
elu_prime (float out)
{
  return out > 0.0f ? 1.0f : out + 1.0f;
}

constexpr int LEARN_ROWS = 16;       /* transitions per block (one warp) */

__global__ void __launch_bounds__ (32)
k_learn_grad (const __grid_constant__ Static st, const LearnArgs a)
{
  extern __shared__ float G[];       /* image of one node's parameter vector: the block's gradient sum */
  const int lane = threadIdx.x;
  const int M = st.M, K = st.K;
  const int P = (int) qnet_floats_per_node (M, K);
  const int chunks = (a.B + LEARN_ROWS - 1) / LEARN_ROWS;
  const int i = blockIdx.x / chunks, row0 = (blockIdx.x % chunks) * LEARN_ROWS;
  if (a.active[i] == 0.0f) return;
  for (int p = lane; p < P; p += 32) G[p] = 0.0f;
  __syncwarp ();
  const int k = st.agent_k[i];
  const QPtr w = q_ptrs (a.W + (size_t) i * P, M, K);
  /* offsets of the layers inside G (the same walk as q_ptrs) */
  float *g_wdst = G, *g_bdst = g_wdst + (size_t) M * QH1, *g_wbuf = g_bdst + QH1, *g_bbuf = g_wbuf + (size_t) K * QH1;
  float *g_w1 = g_bbuf + QH1, *g_b1 = g_w1 + 2 * QH1 * QH2, *g_w2 = g_b1 + QH2, *g_b2 = g_w2 + QH2 * QH2, *g_w3 = g_b2 + QH2;
  float *g_b3 = g_w3 + (size_t) QH2 * K;
  const int ow = 1 + K;
  const int me_under = st.o2u[i];
  float td_sq = 0.0f;
  for (int rr = row0; rr < row0 + LEARN_ROWS && rr < a.B; rr++)
    {
      const size_t ri = (size_t) i * a.B + rr;
      int act = a.act[ri];
      const bool act_ok = act >= 0 && act < k;
      act = act < 0 ? 0 : (act >= k ? k - 1 : act);
      /* ---- TD target ---- */
      float target = a.rew[ri];
      if (!a.target_mode && act_ok)
        {
          const int j_under = st.agent_nbr[(size_t) i * K + act];
          const int j = st.map_overlay[j_under];
          const int kj = st.agent_k[j];
          QRow t;
          q_forward_row (q_ptrs (a.Wt + (size_t) j * P, M, K), M, K, kj, a.nxt + ri * ow, lane, t, st.q_hidden);
          float tq = (lane < kj && (int) st.agent_nbr[(size_t) j * K + lane] != me_under) ? t.q : INFINITY;
          for (int s = 16; s > 0; s >>= 1) tq = fminf (tq, __shfl_xor_sync (0xFFFFFFFFu, tq, s));
          if (tq < INFINITY) target = a.rew[ri] + a.gamma * (1.0f - a.done[ri]) * tq;
        }
      /* ---- forward, loss ---- */
      QRow r;
      q_forward_row (w, M, K, k, a.obs + ri * ow, lane, r, st.q_hidden);
      const float q_sel = __shfl_sync (0xFFFFFFFFu, r.q, act);
      const float td = q_sel - target;
      td_sq += td * td;
      float g = td < -1.0f ? -1.0f : (td > 1.0f ? 1.0f : td);          /* d huber / d td */
      g *= (a.wts ? a.wts[ri] : 1.0f) / (float) a.B;
      /* ---- backward ---- */
      const float d3 = g * elu_prime (q_sel);                           /* at unit `act` of the last layer only */
      g_w3[(size_t) lane * K + act] += r.h2a * d3;
      g_w3[(size_t) (lane + 32) * K + act] += r.h2b * d3;
      if (lane == 0) g_b3[act] += d3;
      const float dz2a = w.w3[(size_t) lane * K + act] * d3 * elu_prime (r.h2a);
      const float dz2b = w.w3[(size_t) (lane + 32) * K + act] * d3 * elu_prime (r.h2b);
      float s0 = 0.0f, s1 = 0.0f;
      float dz1a = dz2a, dz1b = dz2b;                /* one hidden layer: h2 IS h1, the derivative above is already dz1 */
      if (st.q_hidden > 1)
        {
#pragma unroll 4
          for (int u = 0; u < 64; u++)
            {
              const float h = __shfl_sync (0xFFFFFFFFu, u < 32 ? r.h1a : r.h1b, u & 31);     /* input unit u of layer 2 */
              g_w2[u * QH2 + lane] += h * dz2a;
              g_w2[u * QH2 + lane + 32] += h * dz2b;
              const float dz = __shfl_sync (0xFFFFFFFFu, u < 32 ? dz2a : dz2b, u & 31);      /* output unit u of layer 2 */
              s0 = fmaf (w.w2[(size_t) lane * QH2 + u], dz, s0);
              s1 = fmaf (w.w2[(size_t) (lane + 32) * QH2 + u], dz, s1);
            }
          g_b2[lane] += dz2a; g_b2[lane + 32] += dz2b;
          dz1a = s0 * elu_prime (r.h1a); dz1b = s1 * elu_prime (r.h1b);
        }
      s0 = 0.0f; s1 = 0.0f;
#pragma unroll 4
      for (int u = 0; u < 64; u++)
        {
          const float h = __shfl_sync (0xFFFFFFFFu, u < 32 ? r.hd : r.hb, u & 31);
          g_w1[u * QH2 + lane] += h * dz1a;
          g_w1[u * QH2 + lane + 32] += h * dz1b;
          const float dz = __shfl_sync (0xFFFFFFFFu, u < 32 ? dz1a : dz1b, u & 31);
          s0 = fmaf (w.w1[(size_t) lane * QH2 + u], dz, s0);                 /* d / d hd[lane] */
          s1 = fmaf (w.w1[(size_t) (lane + 32) * QH2 + u], dz, s1);          /* d / d hb[lane] */
        }
      g_b1[lane] += dz1a; g_b1[lane + 32] += dz1b;
      const float dzd = s0 * elu_prime (r.hd), dzb = s1 * elu_prime (r.hb);
      g_wdst[(size_t) r.dst * QH1 + lane] += dzd;
      g_bdst[lane] += dzd;
      for (int c = 0; c < k; c++) g_wbuf[(size_t) c * QH1 + lane] += __shfl_sync (0xFFFFFFFFu, r.xn, c) * dzb;
      g_bbuf[lane] += dzb;
      __syncwarp ();
    }
  float *out = a.grad + (size_t) i * P;
  for (int p = lane; p < P; p += 32)
    if (G[p] != 0.0f) atomicAdd (&out[p], G[p]);
  if (lane == 0) atomicAdd (&a.td_sq[i], td_sq);
}

/* Adam, Keras' defaults (learner.py:103); nodes without a batch on any rank keep their state.  bucket = [M][P]
 * gradients summed over the ranks, then M activity counts. */
__global__ void
k_learn_adam (float *W, float *exp_avg, float *exp_avg_sq, const float *steps, const float *bucket, int M, int P, float inv_world,
              float lr, float b1, float b2, float eps)
{
  const int m = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P || bucket[(size_t) M * P + m] <= 0.0f) return;
  const size_t i = (size_t) m * P + p;
  const float g = bucket[i] * inv_world;
  const float t = steps[m] + 1.0f;
  const float mo = b1 * exp_avg[i] + (1.0f - b1) * g;
  const float ve = b2 * exp_avg_sq[i] + (1.0f - b2) * g * g;
  exp_avg[i] = mo; exp_avg_sq[i] = ve;
  const float m_hat = mo / (1.0f - powf (b1, t)), v_hat = ve / (1.0f - powf (b2, t));
  W[i] -= lr * m_hat / (sqrtf (v_hat) + eps);
}

__global__ void
k_learn_bump (float *steps, const float *bucket, int M, int P)
{
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m < M && bucket[(size_t) M * P + m] > 0.0f) steps[m] += 1.0f;
}
