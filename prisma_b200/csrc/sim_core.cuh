/* prisma_b200 — simulator core: one environment's discrete-event step loop.
 *
 * The code in this header is the body of the sm_100a kernels in prisma_b200.cu (one group of G lanes
 * per environment: all lanes scan the pending-timer slots, the group leader executes the event).  It is
 * written against plain pointers so that tests/host_harness can ALSO compile it with g++ and single-step
 * it in the GPU-less build container; that host build is a debugging aid of the test-suite and is never
 * part of libprisma_b200.so.
 *
 * Semantics follow the reference event by event (paths relative to the reference checkout):
 *   scheduler order (ts, uid)                 ns3-gym/src/core/model/scheduler.h:153-169
 *   Send / DropTail enqueue                   prisma/ns3/point-to-point-net-device.cc:594-666,
 *                                             ns3-gym/src/network/utils/queue.h:444-468
 *   TransmitStart / TransmitComplete          prisma/ns3/point-to-point-net-device.cc:272-336
 *   channel propagation                       ns3-gym/src/point-to-point/model/point-to-point-channel.cc:86-107
 *   Receive + statistics                      prisma/ns3/point-to-point-net-device.cc:370-467
 *   agent hook / observation / action         prisma/ns3/packet-routing-gym.cc:150-267,
 *                                             prisma/ns3/data-packet-manager.cc:161-299
 *   pings                                     prisma/ns3/data-packet-manager.cc:349-413,
 *                                             prisma/ns3/ping-forward-packet-manager.cc:93-156,
 *                                             prisma/ns3/ping-back-packet-manager.cc:93-145
 *   small signalling                          prisma/ns3/data-packet-manager.cc:301-347
 *   Poisson sources                           prisma/ns3/poisson-application.cc:265-358, sim.cc:599-632
 *   Forwarder bookkeeping (reward)            prisma/source/forwarder.py:135-161,334-379
 */
#pragma once

#include <stdint.h>
#include <string.h>
#include <math.h>

#include "../../include/prisma_b200.h"

#if defined(__CUDACC__)
#define PRB_HD __host__ __device__ __forceinline__
#define PRB_HDN __host__ __device__
#define PRB_NOINLINE __host__ __device__ __noinline__
#else
#define PRB_HD inline
#define PRB_HDN
#define PRB_NOINLINE
#endif

namespace prb {

constexpr int64_t T_IDLE = INT64_MAX;
constexpr uint16_t NIL = 0xFFFF;
constexpr uint32_t KEY_IDLE = 0xFFFFFFFFu;   /* scan key of an unarmed slot */
constexpr uint32_t KEY_FAR = 0xFFFFFFFEu;    /* saturated: at least 2^32-2 ns after the epoch */
constexpr int NODE_NONE = 255;         /* packed stand-in for lastHop = 1000 / prev_node = -1 */

enum { MODE_RUN = 0, MODE_GYM = 1 };
enum { ST_CONTINUE = 0, ST_PAUSE = 1 };
enum { SLOT_NONE = -1, SLOT_RESUME = -2 };   /* SLOT_RESUME: finish the suspended Receive of gym mode */

/* size classes of packets on the wire */
enum { SZC_DATA_INJ = 0, SZC_DATA_FWD = 1, SZC_PING = 2, SZC_SIGN0 = 3, MAX_SZC = 16 };

/* ---------------------------------------------------------------------------------------------
 * Packet record (lives in the per-env pool in HBM; 40 bytes)
 * ------------------------------------------------------------------------------------------- */
struct Pkt
{
  uint64_t w0;       /* packed: type[0:3) valable[3] rejected[4] first_hop[5] tracked[6] szc[7:11) tunnel[11:15)
                        src[16:24) dst[24:32) last_hop[32:40) next_hop[40:48) ipdst[48:56) prev_node[56:64) */
  int64_t prev_t;    /* time of the previous agent decision on this packet (Agent.temp_obs[...]['time']) */
  int64_t rcv_ts;    /* while on the wire: arrival time and scheduler uid of its Receive event */
  uint32_t rcv_seq;
  uint32_t uid;      /* data: packet uid; small signalling: uid it reports; ping: ping index;
                        big signalling: ordinal c of the segment (NNIndex = c / segs, segIndex = c % segs) */
  uint32_t aux;      /* data: start time, whole s; ping fwd: start, ms; ping back: float bits of the delay */
  uint16_t next;     /* FIFO link (queue or wire) / free list */
  uint16_t pad;
};

#define PRB_FIELD(name, lo, bits)                                                                   \
  PRB_HD int name (const Pkt &p) { return (int) ((p.w0 >> lo) & ((1ull << bits) - 1)); }           \
  PRB_HD void set_##name (Pkt &p, int v)                                                            \
  { p.w0 = (p.w0 & ~(((1ull << bits) - 1) << lo)) | (((uint64_t) v & ((1ull << bits) - 1)) << lo); }
PRB_FIELD (p_type, 0, 3)
PRB_FIELD (p_valable, 3, 1)
PRB_FIELD (p_rejected, 4, 1)
PRB_FIELD (p_first_hop, 5, 1)
PRB_FIELD (p_tracked, 6, 1)
PRB_FIELD (p_szc, 7, 4)
PRB_FIELD (p_tunnel, 11, 4)
PRB_FIELD (p_src, 16, 8)
PRB_FIELD (p_dst, 24, 8)
PRB_FIELD (p_last_hop, 32, 8)
PRB_FIELD (p_next_hop, 40, 8)
PRB_FIELD (p_ipdst, 48, 8)
PRB_FIELD (p_prev_node, 56, 8)
#undef PRB_FIELD

/* ---------------------------------------------------------------------------------------------
 * Per-handle constants and read-only tables (device pointers inside the kernels)
 * ------------------------------------------------------------------------------------------- */
struct Static
{
  int32_t E, N, M, L, LA, F, K, S, P, MA;
  int32_t FP;                          /* flows [0,FP) are Poisson sources, [FP,F) NN-segment sources (big signalling) */
  int32_t big_segs, big_szc;           /* segments per NN (bigSignalingSize / 512), wire size class of a segment */
  int64_t big_start_ns, big_gap_ns;    /* Seconds (AppStartTime) and Seconds (512*8 / (size*8/syncStep)) */
  int32_t NK, NKP;                     /* non-flow timer slots (2*LA + M) and their count padded to 32 */
  int32_t n_szc, n_rc;
  int64_t link_delay_ns, sim_time_ns, ping_interval_ns;
  uint32_t max_out_buffer, ma_size;
  int32_t ping_as_obs, train, opt_agent, inject_mode, policy, mode;
  float loss_penalty;
  uint64_t seed, env_id_base, policy_seed;
  int32_t q_hidden, q_pad_;            /* hidden layers after the concatenation of the Q-networks: 2, or 1 (prb_qnet_set_arch) */
  uint64_t inv1e9_hi, inv1e9_lo;       /* int64x64_t::Invert (1e9), computed on the host */
  int64_t pol_n, tr_n;
  int64_t ev_cap, st_cap;
  /* tables */
  const int16_t *link_from, *link_to;  /* [LA]; access link of node i: from = N+i, to = i */
  const uint8_t *link_rc;              /* [LA] rate class */
  const int64_t *tx_ns;                /* [n_rc * MAX_SZC] */
  const uint16_t *szc_bytes;           /* [MAX_SZC] wire bytes incl. PPP */
  const int16_t *route;                /* [N*N] out link of x towards d */
  const int16_t *map_overlay;          /* [N] */
  const int16_t *o2u;                  /* [M] */
  const uint8_t *agent_k;              /* [M] */
  const int16_t *agent_nbr;            /* [M*K] underlay ids in action order */
  const uint8_t *agent_sign_szc;       /* [M] */
  const int16_t *flow_src, *flow_dst;  /* [F] */
  const double *flow_mean_s;           /* [F] */
  const double *opt_rejected;          /* [N*N] or null */
  const int32_t *pol_data;             /* table [M*M] or action trace [pol_n] */
  const int64_t *tr_t;                 /* injection trace */
  const uint32_t *tr_uid;
  const uint8_t *tr_flags;             /* bit0 valable, bit1 rejected */
  const int32_t *tr_next;              /* next injection of the same flow or -1 */
  const int32_t *flow_first;           /* [F] first trace index or -1 */
  /* state blob layout (byte offsets inside one env's blob) */
  int32_t FG;                          /* groups of 32 flow timers (flow-group minima, scan_flow_groups) */
  int32_t off_fg_ts, off_fg_seq, off_fg_slot;
  int32_t off_scal, off_t_ts, off_sent_mask, off_t_key, off_t_seq, off_q_bytes, off_f_pos, off_ping_index, off_win,
      off_lost, off_n_lost, off_q_pkts, off_q_head, off_q_tail, off_w_head, off_w_tail, off_last_dst, off_n_win,
      off_sent_old;
  int32_t blob_bytes;
};

/* scalars of one environment (first thing in its blob) */
struct Scal
{
  int64_t now;
  int64_t epoch;            /* scan keys are (ts - epoch) in 32 bits */
  int64_t front_ts;         /* scheduler frontier: events with (ts, seq) below it have executed */
  uint64_t pol_rng;
  int64_t pol_pos;
  int64_t ev_n, st_n;
  uint32_t seq, stop_seq;
  uint32_t next_uid;
  uint32_t front_seq, pend_seq;
  int32_t started, finished;
  int32_t pending;          /* gym mode: 1 = a step record is out and its Receive is suspended */
  int32_t pend_link, pend_pkt, pend_m, pend_consult, pend_done, pend_data;
  uint16_t free_head;
  uint16_t pad16;
  int32_t pad32;
  prb_counters c;
};

struct Dyn
{
  Scal *sc;
  int64_t *t_ts;
  uint32_t *t_seq;
  uint32_t *t_key;
  int64_t *fg_ts;      /* [FG] earliest timer of flow group g (slots NK + 32 g ..), its scheduler uid and its slot */
  uint32_t *fg_seq;
  int32_t *fg_slot;
  uint64_t *sent_mask;
  uint32_t *q_bytes, *f_pos, *ping_index, *lost, *n_lost;
  float *win;
  uint16_t *q_pkts, *q_head, *q_tail, *w_head, *w_tail;
  int16_t *last_dst;
  uint8_t *n_win, *sent_old;
  Pkt *pool;
  prb_event *ev_log;
  prb_step_rec *st_log;
};

PRB_HD void
bind_state (const Static &st, char *blob, Dyn &d)
{
  d.sc = (Scal *) (blob + st.off_scal);
  d.t_ts = (int64_t *) (blob + st.off_t_ts);
  d.fg_ts = (int64_t *) (blob + st.off_fg_ts);
  d.fg_seq = (uint32_t *) (blob + st.off_fg_seq);
  d.fg_slot = (int32_t *) (blob + st.off_fg_slot);
  d.sent_mask = (uint64_t *) (blob + st.off_sent_mask);
  d.t_seq = (uint32_t *) (blob + st.off_t_seq);
  d.t_key = (uint32_t *) (blob + st.off_t_key);
  d.q_bytes = (uint32_t *) (blob + st.off_q_bytes);
  d.f_pos = (uint32_t *) (blob + st.off_f_pos);
  d.ping_index = (uint32_t *) (blob + st.off_ping_index);
  d.win = (float *) (blob + st.off_win);
  d.lost = (uint32_t *) (blob + st.off_lost);
  d.n_lost = (uint32_t *) (blob + st.off_n_lost);
  d.q_pkts = (uint16_t *) (blob + st.off_q_pkts);
  d.q_head = (uint16_t *) (blob + st.off_q_head);
  d.q_tail = (uint16_t *) (blob + st.off_q_tail);
  d.w_head = (uint16_t *) (blob + st.off_w_head);
  d.w_tail = (uint16_t *) (blob + st.off_w_tail);
  d.last_dst = (int16_t *) (blob + st.off_last_dst);
  d.n_win = (uint8_t *) (blob + st.off_n_win);
  d.sent_old = (uint8_t *) (blob + st.off_sent_old);
}

/* timer slots: [0,LA) TransmitComplete of link l; [LA,2LA) head-of-wire Receive of link l;
 * [2LA,2LA+M) ping timer of overlay node m; [2LA+M, 2LA+M+F) next injection of flow f.
 * The first NK = 2LA+M slots also carry a 32-bit scan key (t_key) that the group scans per event.  A
 * TransmitComplete whose queue is empty does nothing but mark the device idle, so its slot stays unarmed
 * (key idle) and the event is elided: Send compares its (ts, uid) with the running event's instead, and
 * the slot is armed only once a packet waits behind the one on the wire. */
PRB_HD int slot_txc (const Static &st, int l) { (void) st; return l; }
PRB_HD int slot_rcv (const Static &st, int l) { return st.LA + l; }
PRB_HD int slot_ping (const Static &st, int m) { return 2 * st.LA + m; }
PRB_HD int slot_flow (const Static &st, int f) { return 2 * st.LA + st.M + f; }

/* ---------------------------------------------------------------------------------------------
 * Arithmetic that has to be bit-identical on host and device
 * ------------------------------------------------------------------------------------------- */
PRB_HD uint64_t
umul64hi_ (uint64_t a, uint64_t b)
{
#if defined(__CUDA_ARCH__)
  return __umul64hi (a, b);
#else
  return (uint64_t) (((unsigned __int128) a * b) >> 64);
#endif
}

PRB_HD int
clz64_ (uint64_t x)
{
#if defined(__CUDA_ARCH__)
  return __clzll ((long long) x);
#else
  return x ? __builtin_clzll (x) : 64;
#endif
}

PRB_HD double
u64_to_double_rn (uint64_t x)
{
#if defined(__CUDA_ARCH__)
  return __ull2double_rn (x);
#else
  return (double) x;
#endif
}

PRB_HD double
pow2_ (int e)   /* 2^e, -1022 <= e <= 1023 */
{
  uint64_t b = (uint64_t) (1023 + e) << 52;
  double r;
#if defined(__CUDA_ARCH__)
  r = __longlong_as_double ((long long) b);
#else
  memcpy (&r, &b, 8);
#endif
  return r;
}

/* Time::GetSeconds () for ns >= 0: To(S) multiplies by Invert(1e9) in Q64.64 (nstime.h:505-523,
 * int64x64-128.cc:191-214), GetDouble adds the two words as long double and narrows to double
 * (int64x64-128.h:203-213).  The x87 sum rounds to a 64-bit significand, the narrowing to 53 bits:
 * both roundings are reproduced here with integers. */
/* int64x64_t::Invert (1e9) = floor (2^128 / 1e9) (+0: the round trip gives 1); model.h recomputes it and
 * refuses to build a model if it differs */
constexpr uint64_t INV1E9_HI = 18446744073ull;
constexpr uint64_t INV1E9_LO = 13088917067439035464ull;

#if defined(__CUDACC__)
__host__ __device__ __noinline__
#else
inline
#endif
double
ns_to_seconds_ (int64_t ns)
{
  uint64_t a = (uint64_t) ns;
  /* q = a*inv_hi + ((a*inv_lo) >> 64), a 128-bit Q64.64 value */
  uint64_t lo1 = a * INV1E9_HI, hi1 = umul64hi_ (a, INV1E9_HI);
  uint64_t add = umul64hi_ (a, INV1E9_LO);
  uint64_t q_lo = lo1 + add;
  uint64_t q_hi = hi1 + (q_lo < lo1 ? 1 : 0);
  if (q_hi == 0)
    return u64_to_double_rn (q_lo) * pow2_ (-64);
  int lz = clz64_ (q_hi);
  int s = 64 - lz;                       /* bits to shed so that 64 significant bits remain */
  uint64_t m = (lz ? (q_hi << lz) : q_hi) | (lz ? (q_lo >> s) : 0);
  if (lz == 0) m = q_hi;
  uint64_t rem = (s == 64) ? q_lo : (q_lo & ((1ull << s) - 1));
  uint64_t half = 1ull << (s - 1);
  if (rem > half || (rem == half && (m & 1)))
    {
      m += 1;
      if (m == 0)
        {
          m = 1ull << 63;
          s += 1;
        }
    }
  return u64_to_double_rn (m) * pow2_ (s - 64);
}

PRB_HD double
ns_to_seconds (const Static &st, int64_t ns)
{
  (void) st;
  return ns_to_seconds_ (ns);
}

PRB_HD double dmul_ (double a, double b)
{
#if defined(__CUDA_ARCH__)
  return __dmul_rn (a, b);
#else
  return a * b;
#endif
}
PRB_HD double dadd_ (double a, double b)
{
#if defined(__CUDA_ARCH__)
  return __dadd_rn (a, b);
#else
  return a + b;
#endif
}
PRB_HD double dsub_ (double a, double b)
{
#if defined(__CUDA_ARCH__)
  return __dsub_rn (a, b);
#else
  return a - b;
#endif
}
PRB_HD double ddiv_ (double a, double b)
{
#if defined(__CUDA_ARCH__)
  return __ddiv_rn (a, b);
#else
  return a / b;
#endif
}
PRB_HD double dfma_ (double a, double b, double c)
{
#if defined(__CUDA_ARCH__)
  return __fma_rn (a, b, c);
#else
  return fma (a, b, c);
#endif
}
PRB_HD float d2f_ (double a)
{
#if defined(__CUDA_ARCH__)
  return __double2float_rn (a);
#else
  return (float) a;
#endif
}
PRB_HD float fadd_ (float a, float b)
{
#if defined(__CUDA_ARCH__)
  return __fadd_rn (a, b);
#else
  return a + b;
#endif
}

/* Philox-4x32-10 (Salmon et al., SC'11) */
#ifndef PRB_PHILOX_UNROLL
#define PRB_PHILOX_UNROLL 10
#endif
constexpr int PHILOX_UNROLL = PRB_PHILOX_UNROLL;
PRB_HD void
philox4x32 (uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4])
{
#pragma unroll PHILOX_UNROLL
  for (int r = 0; r < 10; r++)
    {
      uint64_t p0 = (uint64_t) 0xD2511F53u * c0;
      uint64_t p1 = (uint64_t) 0xCD9E8D57u * c2;
      uint32_t n0 = (uint32_t) (p1 >> 32) ^ c1 ^ k0;
      uint32_t n1 = (uint32_t) p1;
      uint32_t n2 = (uint32_t) (p0 >> 32) ^ c3 ^ k1;
      uint32_t n3 = (uint32_t) p0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      k0 += 0x9E3779B9u;
      k1 += 0xBB67AE85u;
    }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* u in (0,1) from 52 random bits */
PRB_HD double
u01_53 (uint32_t a, uint32_t b)
{
  uint64_t x = (((uint64_t) a << 32) | b) >> 12;
  return dmul_ (dadd_ (u64_to_double_rn (x), 0.5), 1.0 / 4503599627370496.0);
}

/* -log(u), u in (0,1), from IEEE +,-,*,/ and fma only (the synthetic-traffic specification:
 * DESIGN.md "Synthetic traffic"; ns-3's own libm call is matched by trace replay instead) */
PRB_HD double
neglog (double u)
{
  uint64_t bits;
#if defined(__CUDA_ARCH__)
  bits = (uint64_t) __double_as_longlong (u);
#else
  memcpy (&bits, &u, 8);
#endif
  int e = (int) ((bits >> 52) & 0x7FF) - 1023;
  uint64_t mb = (bits & 0xFFFFFFFFFFFFFull) | 0x3FF0000000000000ull;
  double m;
#if defined(__CUDA_ARCH__)
  m = __longlong_as_double ((long long) mb);
#else
  memcpy (&m, &mb, 8);
#endif
  if (m > 1.4142135623730951)
    {
      m = dmul_ (m, 0.5);
      e += 1;
    }
  double s = ddiv_ (dsub_ (m, 1.0), dadd_ (m, 1.0));
  double z = dmul_ (s, s);
  double p = 1.0 / 27.0;
  p = dfma_ (p, z, 1.0 / 25.0);
  p = dfma_ (p, z, 1.0 / 23.0);
  p = dfma_ (p, z, 1.0 / 21.0);
  p = dfma_ (p, z, 1.0 / 19.0);
  p = dfma_ (p, z, 1.0 / 17.0);
  p = dfma_ (p, z, 1.0 / 15.0);
  p = dfma_ (p, z, 1.0 / 13.0);
  p = dfma_ (p, z, 1.0 / 11.0);
  p = dfma_ (p, z, 1.0 / 9.0);
  p = dfma_ (p, z, 1.0 / 7.0);
  p = dfma_ (p, z, 1.0 / 5.0);
  p = dfma_ (p, z, 1.0 / 3.0);
  p = dfma_ (p, z, 1.0);
  double lm = dmul_ (dmul_ (2.0, s), p);
  double r = dfma_ ((double) e, 0.6931471805599453, lm);
  return -r;
}

/* ---------------------------------------------------------------------------------------------
 * The simulator of one environment.  Every method is executed by the group leader only.
 * ------------------------------------------------------------------------------------------- */
template <bool LOG, bool GYM>
struct Sim
{
  /* the Gym step record is assembled only when somebody reads it: the parity log or prb_step */
  static constexpr bool REC = LOG || GYM;
  const Static &st;
  Dyn &d;
  Scal *sc;
  int env;
  uint32_t cur_seq;      /* scheduler uid of the running event */
  /* Sends requested by the running event are executed after its handler, in order, by ONE inlined copy of
   * Send/TransmitStart (flush_sends): the first waits in registers, further ones are chained through the
   * pool (link id parked in rcv_seq).  post_kind: what the event schedules after its sends. */
  int ps_n, ps_l0, ps_pi0;
  Pkt ps_q0;
  uint16_t ps_head, ps_tail;
  int post_kind, post_arg;
  prb_step_rec *rec;     /* where the current Gym step is assembled */

  PRB_HD Sim (const Static &st_, Dyn &d_, int env_, prb_step_rec *rec_) : st (st_), d (d_), sc (d_.sc), env (env_), cur_seq (0), ps_n (0), ps_l0 (0), ps_pi0 (0), ps_head (NIL), ps_tail (NIL), post_kind (0), post_arg (0), rec (rec_) {}

  /* ---- timers ---- */
  PRB_HD uint32_t
  mk_key (int64_t ts) const
  {
    uint64_t dlt = (uint64_t) (ts - sc->epoch);
    return dlt < (uint64_t) KEY_FAR ? (uint32_t) dlt : KEY_FAR;
  }
  PRB_HD void
  set_timer (int slot, int64_t ts)
  {
    d.t_ts[slot] = ts;
    d.t_seq[slot] = sc->seq++;
    if (slot < st.NK) d.t_key[slot] = mk_key (ts);
  }
  PRB_HD void
  put_timer (int slot, int64_t ts, uint32_t seq)
  {
    d.t_ts[slot] = ts;
    d.t_seq[slot] = seq;
    if (slot < st.NK) d.t_key[slot] = mk_key (ts);
  }
  PRB_HD void
  clear_timer (int slot)
  {
    d.t_ts[slot] = T_IDLE;
    if (slot < st.NK) d.t_key[slot] = KEY_IDLE;
  }
  /* new epoch: every armed key is rewritten from the 64-bit truth */
  PRB_HD void
  rebase (int64_t epoch)
  {
    sc->epoch = epoch;
    for (int s2 = 0; s2 < st.NK; s2++)
      if (d.t_key[s2] != KEY_IDLE) d.t_key[s2] = mk_key (d.t_ts[s2]);
  }

  /* ---- packet pool ---- */
  PRB_HD int
  alloc ()
  {
    uint16_t h = sc->free_head;
    if (h == NIL)
      {
        sc->c.errors |= PRB_ERR_POOL;
        return -1;
      }
    sc->free_head = d.pool[h].next;
    return h;
  }
  PRB_HD void
  release (int i)
  {
    d.pool[i].next = sc->free_head;
    sc->free_head = (uint16_t) i;
  }

  /* ---- logs ---- */
  PRB_HD void
  log_event (int kind, int node, int link, const Pkt &p, int qbytes)
  {
    if (!LOG) return;
    if (sc->ev_n >= st.ev_cap)
      {
        if (st.ev_cap > 0) sc->c.errors |= PRB_ERR_LOG;
        return;
      }
    prb_event *e = &d.ev_log[sc->ev_n++];
    int data = p_type (p) == PRB_DATA;
    e->t_ns = sc->now; e->kind = kind; e->node = node; e->link = link; e->pad_ = 0;
    e->uid = data ? (int64_t) p.uid : -1;
    e->size = st.szc_bytes[p_szc (p)]; e->type = p_type (p); e->src = p_src (p); e->dst = p_dst (p);
    e->last_hop = p_last_hop (p) == NODE_NONE ? PRB_LASTHOP_NONE : p_last_hop (p);
    e->next_hop = p_next_hop (p); e->valable = data ? p_valable (p) : -1;
    e->qbytes = qbytes;
  }

  PRB_HD void
  push_f (float *sum, int64_t *n, float v)
  {
    /* ComputeStats re-sums a float vector front to back (compute-stats-v2.cc:87-97): same value as a
     * running float sum in push order */
    *sum = fadd_ (*sum, v);
    *n += 1;
  }

  /* MacTxDrop fan-out — data-packet-manager.cc:87-110 */
  PRB_HD void
  mac_tx_drop (int link, const Pkt &p)
  {
    if (link >= st.L) return;
    if (p_type (p) != PRB_DATA || !p_valable (p)) return;
    int lh = p_last_hop (p);
    if (lh == NODE_NONE) return;
    int m = st.map_overlay[lh];
    if (m < 0 || p_dst (p) == lh) return;
    uint32_t n = d.n_lost[m];
    d.lost[m * PRB_MAX_LOST + (n % PRB_MAX_LOST)] = p.uid;
    d.n_lost[m] = n + 1;
  }

  /* PointToPointNetDevice::TransmitStart + PointToPointChannel::TransmitStart */
  PRB_HD void
  transmit_start (int l, int pi, Pkt &p)
  {
    if (l < st.L) log_event (PRB_EV_TXB, st.link_from[l], l, p, (int) d.q_bytes[l]);
    int64_t tx = st.tx_ns[st.link_rc[l] * MAX_SZC + p_szc (p)];
    int64_t delay = l < st.L ? st.link_delay_ns : 0;
    /* TransmitComplete: scheduled (it takes a scheduler uid) but armed only while a packet waits */
    d.t_ts[l] = sc->now + tx;
    d.t_seq[l] = sc->seq++;
    d.t_key[l] = d.q_head[l] != NIL ? mk_key (sc->now + tx) : KEY_IDLE;
    p.rcv_ts = sc->now + tx + delay;
    p.rcv_seq = sc->seq++;
    p.next = NIL;
    d.pool[pi] = p;
    uint16_t t = d.w_tail[l];
    if (t == NIL)
      {
        d.w_head[l] = (uint16_t) pi;
        put_timer (slot_rcv (st, l), p.rcv_ts, p.rcv_seq);
      }
    else
      d.pool[t].next = (uint16_t) pi;
    d.w_tail[l] = (uint16_t) pi;
  }

  PRB_HD void
  count_drop (const Pkt &p, bool overlay_side)
  {
    if (overlay_side)
      {
        sc->c.overlay_lost++;
        push_f (&sc->c.sum_cost, &sc->c.n_cost, st.loss_penalty);
      }
    else
      {
        push_f (&sc->c.sum_underlay_cost, &sc->c.n_underlay_cost, st.loss_penalty);
        sc->c.underlay_lost++;
      }
    (void) p;
  }

  /* PointToPointNetDevice::Send + Queue::DoEnqueue.  p is the register copy of pool[pi]; takes ownership. */
  PRB_HD void
  dev_send (int l, int pi, Pkt &p)
  {
    int node = st.link_from[l];
    int data = p_type (p) == PRB_DATA;
    if (data && p_rejected (p) && node < 23)
      {
        count_drop (p, p_valable (p));
        mac_tx_drop (l, p);
        release (pi);
        return;
      }
    if (l >= st.L) log_event (PRB_EV_INJ, node, l, p, -1);
    uint32_t size = st.szc_bytes[p_szc (p)];
    uint32_t qb = d.q_bytes[l];
    uint32_t qp = d.q_pkts[l];
    bool fits = l < st.L ? (qb + size <= st.max_out_buffer) : (qp + 1 <= 1000u);
    if (!fits)
      {
        if (l < st.L) log_event (PRB_EV_DRP, node, l, p, (int) qb);
        if (data)
          {
            count_drop (p, p_valable (p) && p_dst (p) != node);
            mac_tx_drop (l, p);
          }
        release (pi);
        return;
      }
    if (l < st.L) log_event (PRB_EV_ENQ, node, l, p, (int) (qb + size));
    int64_t tc = d.t_ts[l];
    bool busy = tc != T_IDLE;
    if (busy && d.t_key[l] == KEY_IDLE && (tc < sc->now || (tc == sc->now && d.t_seq[l] < cur_seq)))
      {
        /* the elided TransmitComplete lies before the running event in scheduler order: it has happened */
        sc->c.events++;
        d.t_ts[l] = T_IDLE;
        busy = false;
      }
    if (!busy)
      {
        /* READY: the packet goes through the (empty) queue straight to the wire */
        transmit_start (l, pi, p);
        return;
      }
    if (d.t_key[l] == KEY_IDLE) d.t_key[l] = mk_key (tc);   /* a packet waits: the completion becomes a real event */
    p.next = NIL;
    d.pool[pi] = p;
    uint16_t t = d.q_tail[l];
    if (t == NIL) d.q_head[l] = (uint16_t) pi; else d.pool[t].next = (uint16_t) pi;
    d.q_tail[l] = (uint16_t) pi;
    d.q_bytes[l] = qb + size;
    d.q_pkts[l] = (uint16_t) (qp + 1);
  }

  PRB_HD void
  post_send (int l, int pi, const Pkt &p)
  {
    if (ps_n == 0)
      {
        ps_l0 = l; ps_pi0 = pi; ps_q0 = p;
      }
    else
      {
        Pkt q = p;
        q.rcv_seq = (uint32_t) l;
        q.next = NIL;
        d.pool[pi] = q;
        if (ps_tail == NIL) ps_head = (uint16_t) pi; else d.pool[ps_tail].next = (uint16_t) pi;
        ps_tail = (uint16_t) pi;
      }
    ps_n++;
  }

  PRB_HD void
  flush_sends ()
  {
    while (ps_n > 0)
      {
        int l = ps_l0, pi = ps_pi0;
        Pkt p = ps_q0;
        ps_n--;
        if (ps_n > 0)
          {
            /* next request comes out of the chain */
            ps_pi0 = ps_head;
            ps_q0 = d.pool[ps_head];
            ps_l0 = (int) ps_q0.rcv_seq;
            ps_head = ps_q0.next;
            if (ps_head == NIL) ps_tail = NIL;
          }
        dev_send (l, pi, p);
      }
  }

  /* PointToPointNetDevice::TransmitComplete */
  PRB_HD void
  transmit_complete (int l)
  {
    clear_timer (slot_txc (st, l));
    uint16_t h = d.q_head[l];
    if (h == NIL) return;
    Pkt p = d.pool[h];
    d.q_head[l] = p.next;
    if (p.next == NIL) d.q_tail[l] = NIL;
    d.q_bytes[l] -= st.szc_bytes[p_szc (p)];
    d.q_pkts[l] -= 1;
    transmit_start (l, h, p);
  }

  /* ---- agents ---- */
  PRB_HD int64_t
  ping_start_ms (uint32_t idx) const
  {
    /* ping idx fires at (idx+1)*interval; Time::GetMilliSeconds truncates */
    return ((int64_t) (idx + 1) * st.ping_interval_ns) / 1000000;
  }

  /* PingBackPacketManager::getMaxTimePingForwardPacketSent */
  PRB_HD float
  ping_oldest_age (int m, int t)
  {
    int i = m * st.K + t;
    if (d.sent_old[i]) return d2f_ (2.60);   /* an unacknowledged ping older than the window: the cap applies */
    uint64_t mask = d.sent_mask[i];
    if (mask == 0) return 0.0f;
    int j = 63 - clz64_ (mask);
    uint32_t idx = d.ping_index[m] - 1 - (uint32_t) j;
    double v = dsub_ (ns_to_seconds (st, sc->now), dmul_ ((double) ping_start_ms (idx), 0.001));
    return d2f_ (v < 2.60 ? v : 2.60);
  }

  /* DataPacketManager::getObservation */
  PRB_HD int
  build_obs (int m, const Pkt &p, int32_t *obs)
  {
    int node = st.o2u[m];
    int k = st.agent_k[m];
    obs[0] = st.map_overlay[p_dst (p)];
    for (int i = 0; i < k; i++)
      {
        int out = st.route[node * st.N + st.agent_nbr[m * st.K + i]];
        uint32_t value;
        if (!st.ping_as_obs)
          value = d.q_bytes[out];
        else
          {
            double avg = 0.0;
            int nw = d.n_win[m * st.K + i];
            if (nw > 0)
              {
                double sum = 0.0;
                for (int j = 0; j < nw; j++) sum = dadd_ (sum, (double) d.win[(m * st.K + i) * st.MA + j]);
                avg = ddiv_ (sum, (double) nw);
              }
            double age = (double) ping_oldest_age (m, i);
            value = (uint32_t) dmul_ (1000.0, (avg > age ? avg : age));
          }
        obs[1 + i] = (int32_t) value;
      }
    return 1 + k;
  }

  PRB_HD int32_t
  policy_action (int m, const int32_t *obs, int k)
  {
    switch (st.policy)
      {
      case PRB_POLICY_TABLE:
        return st.pol_data[m * st.M + obs[0]];
      case PRB_POLICY_TRACE:
        if (sc->pol_pos >= st.pol_n)
          {
            sc->c.errors |= PRB_ERR_TRACE;
            return 0;
          }
        return st.pol_data[sc->pol_pos++];
      case PRB_POLICY_RANDOM:
        {
          uint64_t x = sc->pol_rng;
          x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
          sc->pol_rng = x;
          return (int32_t) ((x * 2685821657736338717ULL) % (uint64_t) k);
        }
      default:
        return 0;
      }
  }

  /* DataPacketManager::sendSmallSignalingPacket */
  PRB_HD void
  send_small_signaling (int m, const Pkt &p, int in_link)
  {
    int node = st.o2u[m];
    if (p_src (p) == node) return;
    int qi = alloc ();
    if (qi < 0) return;
    Pkt q;
    q.w0 = 0; q.prev_t = 0; q.rcv_ts = 0; q.rcv_seq = 0; q.pad = 0; q.next = NIL;
    set_p_type (q, PRB_SMALL_SIGN);
    int lh = p_last_hop (p);
    set_p_dst (q, lh); set_p_next_hop (q, lh); set_p_last_hop (q, node); set_p_src (q, node); set_p_ipdst (q, lh);
    set_p_prev_node (q, NODE_NONE);
    set_p_szc (q, st.agent_sign_szc[m]);
    q.uid = p.uid; q.aux = 0;
    post_send (in_link ^ 1, qi, q);
  }

  /* PingForwardPacketManager::receivePacket / sendPingBackPacket */
  PRB_HD void
  ping_forward_rx (int m, const Pkt &p, int in_link)
  {
    int node = st.o2u[m];
    float delay = d2f_ (dsub_ (ns_to_seconds (st, sc->now), dmul_ ((double) p.aux, 0.001)));
    int qi = alloc ();
    if (qi < 0) return;
    Pkt q;
    q.w0 = 0; q.prev_t = 0; q.rcv_ts = 0; q.rcv_seq = 0; q.pad = 0; q.next = NIL;
    set_p_type (q, PRB_PING_BACK);
    int lh = p_last_hop (p);
    set_p_dst (q, lh); set_p_next_hop (q, lh); set_p_last_hop (q, node); set_p_src (q, node); set_p_ipdst (q, lh);
    set_p_prev_node (q, NODE_NONE);
    set_p_szc (q, SZC_PING);
    set_p_tunnel (q, p_tunnel (p));
    q.uid = p.uid;
#if defined(__CUDA_ARCH__)
    q.aux = __float_as_uint (delay);
#else
    memcpy (&q.aux, &delay, 4);
#endif
    post_send (in_link ^ 1, qi, q);
  }

  /* PingBackPacketManager::receivePacket */
  PRB_HD void
  ping_back_rx (int m, const Pkt &p)
  {
    int t = p_tunnel (p);
    int i = m * st.K + t;
    if (t >= st.K) return;
    uint32_t j = d.ping_index[m] - 1 - p.uid;
    if (j < 64)
      d.sent_mask[i] &= ~(1ull << j);
    int nw = d.n_win[i];
    float *w = &d.win[i * st.MA];
    if ((uint32_t) nw >= st.ma_size && nw > 0)
      {
        for (int a = 0; a + 1 < nw; a++) w[a] = w[a + 1];
        nw--;
      }
    float v;
#if defined(__CUDA_ARCH__)
    v = __uint_as_float (p.aux);
#else
    memcpy (&v, &p.aux, 4);
#endif
    w[nw++] = v;
    d.n_win[i] = (uint8_t) nw;
  }

  /* DataPacketManager::sendPingPackets / sendPingForwardPacket + addSentPingForwardPacket */
  PRB_HD void
  ping_timer (int m)
  {
    int node = st.o2u[m];
    int k = st.agent_k[m];
    uint32_t idx = d.ping_index[m];
    uint32_t ms = (uint32_t) (sc->now / 1000000);
    for (int t = 0; t < k; t++)
      {
        int i = m * st.K + t;
        uint64_t mask = d.sent_mask[i];
        if (mask >> 63)
          {
            d.sent_old[i] = 1;
            /* the window must cover the 2.60 s cap, else the exact age of the shed ping still matters */
            if (64 * st.ping_interval_ns < 2600000000ll) sc->c.errors |= PRB_ERR_PING_WINDOW;
          }
        d.sent_mask[i] = (mask << 1) | 1ull;
      }
    d.ping_index[m] = idx + 1;
    for (int t = 0; t < k; t++)
      {
        int qi = alloc ();
        if (qi < 0) break;
        int nb = st.agent_nbr[m * st.K + t];
        Pkt q;
        q.w0 = 0; q.prev_t = 0; q.rcv_ts = 0; q.rcv_seq = 0; q.pad = 0; q.next = NIL;
        set_p_type (q, PRB_PING_FWD);
        set_p_dst (q, nb); set_p_next_hop (q, nb); set_p_last_hop (q, node); set_p_src (q, node); set_p_ipdst (q, nb);
        set_p_prev_node (q, NODE_NONE);
        set_p_szc (q, SZC_PING);
        set_p_tunnel (q, t);
        q.uid = idx; q.aux = ms;
        post_send (st.route[node * st.N + nb], qi, q);
      }
    post_kind = 2; post_arg = m;     /* the timer is rescheduled after the sends (data-packet-manager.cc:349-357) */
  }

  /* ---- Receive: part A (up to the agent's observation) ---- */
  PRB_HD int
  receive (int l, int pi)
  {
    Pkt p = d.pool[pi];
    int x = st.link_to[l];
    bool access = l >= st.L;
    int type = p_type (p);
    if (type == PRB_DATA && p_rejected (p) && x < 23)
      {
        mac_tx_drop (access ? l : (l ^ 1), p);
        release (pi);
        return ST_CONTINUE;
      }
    log_event (access ? PRB_EV_ARR : PRB_EV_RCV, x, l, p, -1);
    if (!access)
      {
        sc->c.packet_hops++;
        if (type == PRB_DATA) sc->c.data_hops++;
      }
    int m = st.map_overlay[x];
    bool notify = false;
    if (m >= 0)
      {
        switch (type)
          {
          case PRB_DATA:
            d.last_dst[m] = (int16_t) p_dst (p);
            if (p_next_hop (p) == x && p_valable (p))
              {
                notify = true;
                fill_data_step (m, x, p);
              }
            break;
          case PRB_BIG_SIGN:
            if (p_src (p) != x && p_dst (p) == x && p_next_hop (p) == x)
              {
                notify = true;
                fill_control_step (m, x, p);
              }
            break;
          case PRB_SMALL_SIGN:
            if (p_dst (p) == x && p_next_hop (p) == x)
              {
                notify = true;
                fill_control_step (m, x, p);
              }
            break;
          case PRB_PING_FWD:
            ping_forward_rx (m, p, l);
            break;
          case PRB_PING_BACK:
            ping_back_rx (m, p);
            break;
          }
      }
    if (!notify)
      {
        receive_tail (l, pi, p);
        return ST_CONTINUE;
      }
    sc->c.env_steps++;
    int data = type == PRB_DATA;
    int done = data ? (p_dst (p) == x) : (d.last_dst[m] == x);
    int consult = data && !done;
    if (GYM)
      {
        rec->need_action = consult;
        rec->action = -1;
        rec->valid = 1;
        sc->pending = 1;
        sc->pend_link = l; sc->pend_pkt = pi; sc->pend_m = m; sc->pend_consult = consult;
        sc->pend_done = done; sc->pend_data = data;
        sc->pend_seq = cur_seq;
        sc->front_ts = sc->now; sc->front_seq = cur_seq;
        return ST_PAUSE;
      }
    int32_t action = 0;
    if (consult)
      {
        if (REC)
          action = policy_action (m, rec->obs, st.agent_k[m]);
        else
          {
            /* fused run without a reader of the record: the table agent needs obs[0] only */
            int32_t obs0 = st.map_overlay[p_dst (p)];
            action = policy_action (m, &obs0, st.agent_k[m]);
          }
      }
    if (REC)
      {
        rec->action = consult ? action : -1;
        rec->need_action = 0;
        rec->valid = 1;
        if (LOG && st.st_cap > 0)
          {
            if (sc->st_n < st.st_cap) d.st_log[sc->st_n++] = *rec; else sc->c.errors |= PRB_ERR_LOG;
          }
      }
    finish_step (l, pi, p, m, data, consult, done, action);
    return ST_CONTINUE;
  }

  /* Gym mode: the action of the suspended step has arrived */
  PRB_HD void
  resume (int32_t action)
  {
    int l = sc->pend_link, pi = sc->pend_pkt;
    Pkt p = d.pool[pi];
    sc->pending = 0;
    cur_seq = sc->pend_seq;
    finish_step (l, pi, p, sc->pend_m, sc->pend_data, sc->pend_consult, sc->pend_done, sc->pend_consult ? action : 0);
  }

  /* OpenGymInterface::NotifyCurrentState order: observation etc. first, then ExecuteActions */
  PRB_HD void
  fill_data_step (int m, int x, const Pkt &p)
  {
    if (!REC)
      {
        d.n_lost[m] = 0;      /* the lost list is flushed by every data step (data-packet-manager.cc:229-237) */
        return;
      }
    prb_step_rec *r = rec;
    r->t_ns = sc->now; r->node = x; r->overlay = m; r->uid = (int64_t) p.uid; r->ptype = PRB_DATA;
    r->nobs = build_obs (m, p, r->obs);
    for (int i = r->nobs; i < 1 + PRB_MAX_K; i++) r->obs[i] = 0;
    r->done = p_dst (p) == x;
    r->e2e_delay = d2f_ (dsub_ (ns_to_seconds (st, sc->now), (double) p.aux));
    /* DataPacketManager::getInfo flushes the lost list back to front */
    uint32_t nl = d.n_lost[m];
    r->n_lost = (int32_t) nl;
    for (int i = 0; i < PRB_MAX_LOST; i++)
      r->lost_uid[i] = (uint32_t) i < nl ? (int64_t) d.lost[m * PRB_MAX_LOST + ((nl - 1 - i) % PRB_MAX_LOST)] : 0;
    d.n_lost[m] = 0;
    r->is_new = !p_tracked (p);
    r->prev_node = -1;
    r->hop_time = 0.0;
    if (p_tracked (p))
      {
        r->prev_node = p_prev_node (p);
        r->hop_time = dmul_ ((double) (sc->now - p.prev_t), 1e-9);
      }
    r->src = p_src (p);
  }

  PRB_HD void
  fill_control_step (int m, int x, const Pkt &p)
  {
    if (!REC) return;
    prb_step_rec *r = rec;
    r->t_ns = sc->now; r->node = x; r->overlay = m;
    r->uid = p_type (p) == PRB_SMALL_SIGN ? (int64_t) p.uid : -1;
    r->ptype = p_type (p); r->nobs = 1; r->obs[0] = 1000;
    for (int i = 1; i < 1 + PRB_MAX_K; i++) r->obs[i] = 0;
    r->action = -1; r->prev_node = -1; r->is_new = 0; r->hop_time = 0.0; r->e2e_delay = 0.0f; r->n_lost = 0;
    if (p_type (p) == PRB_BIG_SIGN)
      {
        /* info fields 18-20: NN Index, segment Index, NodeId Signaled (big-signaling-packet-manager.cc:104-121) */
        uint32_t segs = (uint32_t) st.big_segs;
        r->uid = (int64_t) (((uint64_t) (p.uid / segs) << 32) | (uint64_t) (p.uid % segs));
        r->prev_node = p_src (p);
      }
    for (int i = 0; i < PRB_MAX_LOST; i++) r->lost_uid[i] = 0;
    /* GetGameOver asks the DATA manager, whose m_destination is the last data packet seen here */
    r->done = d.last_dst[m] == x;
    r->src = p_src (p);
  }

  /* PacketRoutingEnv::ExecuteActions + DataPacketManager::sendPacket, then the rest of Receive */
  PRB_HD void
  finish_step (int l, int pi, Pkt &p, int m, int data, int consult, int done, int32_t action)
  {
    if (data)
      {
        int node = st.o2u[m];
        if (consult) sc->c.decisions++;
        if (st.train) send_small_signaling (m, p, l);
        if (!done)
          {
            int k = st.agent_k[m];
            if (action >= 0 && action < k)
              {
                int qi = alloc ();
                if (qi >= 0)
                  {
                    Pkt q = p;
                    int nb = st.agent_nbr[m * st.K + action];
                    set_p_last_hop (q, node); set_p_next_hop (q, nb); set_p_ipdst (q, nb);
                    set_p_first_hop (q, 0);
                    set_p_szc (q, SZC_DATA_FWD);
                    set_p_tracked (q, 1); set_p_prev_node (q, node);
                    q.prev_t = sc->now;
                    post_send (st.route[node * st.N + nb], qi, q);
                  }
              }
            else
              sc->c.vanished++;
          }
      }
    receive_tail (l, pi, p);
  }

  /* the IP forward inside m_rxCallback and the statistics after it */
  PRB_HD void
  receive_tail (int l, int pi, Pkt &p)
  {
    int x = st.link_to[l];
    int type = p_type (p);
    int data = type == PRB_DATA;
    bool forwarded = false;
    /* statistics use the fields of the packet as it arrived; the forward moves the record */
    Pkt o = p;
    if (p_first_hop (p))
      {
        /* addressed to its own switch: local delivery */
      }
    else if (p_ipdst (p) == x)
      {
        if (data && p_last_hop (p) == NODE_NONE && !p_valable (p) && !(p_rejected (p) && st.N + x < 23))
          sc->c.underlay_injected++;    /* the traffic node's own Receive counts it once more */
      }
    else
      {
        forwarded = true;
        post_send (st.route[x * st.N + p_ipdst (p)], pi, p);
      }
    if (data && p_dst (o) == x)
      {
        float dly = d2f_ (dsub_ (ns_to_seconds (st, sc->now), (double) o.aux));
        if (p_valable (o))
          {
            if (p_next_hop (o) == p_dst (o))
              {
                sc->c.overlay_arrived++;
                push_f (&sc->c.sum_cost, &sc->c.n_cost, dly);
                push_f (&sc->c.sum_e2e, &sc->c.n_e2e, dly);
              }
          }
        else
          {
            sc->c.underlay_arrived++;
            push_f (&sc->c.sum_underlay_cost, &sc->c.n_underlay_cost, dly);
            push_f (&sc->c.sum_underlay_e2e, &sc->c.n_underlay_e2e, dly);
          }
      }
    if (type > 0 && p_dst (o) == x)
      sc->c.bytes_signaling += st.szc_bytes[p_szc (o)] - 2;
    if (data && p_last_hop (o) == NODE_NONE)
      {
        if (p_valable (o))
          {
            sc->c.overlay_injected++;
            sc->c.bytes_data += st.szc_bytes[p_szc (o)] - 2;
          }
        else
          sc->c.underlay_injected++;
      }
    if (!forwarded) release (pi);
  }

  /* ---- sources ---- */
  PRB_HD void
  schedule_flow (int f, bool first)
  {
    int slot = slot_flow (st, f);
    if (st.inject_mode == PRB_INJECT_TRACE)
      {
        int32_t i = first ? st.flow_first[f] : st.tr_next[(int32_t) d.f_pos[f]];
        d.f_pos[f] = (uint32_t) i;
        if (i >= 0) set_timer (slot, st.tr_t[i]); else clear_timer (slot);
        return;
      }
    if (f >= st.FP)
      {
        /* BigSignalingGeneratorApplication::ScheduleNextTx — big-signaling-application.cc:236-258: a fixed
         * gap after StartApplication at AppStartTime (sim.cc:644); f_pos = ordinal of the pending segment */
        d.f_pos[f] = first ? 1u : d.f_pos[f] + 1u;
        set_timer (slot, (first ? st.big_start_ns : sc->now) + st.big_gap_ns);
        return;
      }
    uint64_t eid = st.env_id_base + (uint64_t) env;
    uint32_t k0 = (uint32_t) st.seed, k1 = (uint32_t) (st.seed >> 32);
    uint32_t fkey = (uint32_t) (st.flow_src[f] * st.N + st.flow_dst[f]);
    uint32_t r[4];
    int64_t base = sc->now;
    uint32_t nd = d.f_pos[f];
    if (first)
      {
        philox4x32 ((uint32_t) eid, (uint32_t) (eid >> 32), fkey, 0u, k0, k1, r);
        double u0 = u01_53 (r[0], r[1]);
        base = 100000 + (int64_t) dmul_ (u0, 1e9);
        nd = 1;
      }
    philox4x32 ((uint32_t) eid, (uint32_t) (eid >> 32), fkey, nd, k0, k1, r);
    d.f_pos[f] = nd + 1;
    double u = u01_53 (r[0], r[1]);
    double delay = dmul_ (st.flow_mean_s[f], neglog (u));
    int64_t t = base + (int64_t) dmul_ (delay, 1e9);
    set_timer (slot, t);
  }

  /* PoissonGeneratorApplication::SendPacket */
  PRB_HD void
  inject (int f)
  {
    int src = st.flow_src[f], dst = st.flow_dst[f];
    int pi = alloc ();
    Pkt p;
    p.w0 = 0; p.prev_t = 0; p.rcv_ts = 0; p.rcv_seq = 0; p.pad = 0; p.next = NIL;
    if (f >= st.FP)
      {
        /* BigSignalingGeneratorApplication::SendPacket — big-signaling-application.cc:261-306 */
        set_p_type (p, PRB_BIG_SIGN);
        set_p_dst (p, dst); set_p_src (p, src); set_p_next_hop (p, dst); set_p_last_hop (p, NODE_NONE);
        set_p_prev_node (p, NODE_NONE);
        set_p_szc (p, st.big_szc);
        set_p_ipdst (p, dst);
        p.aux = 0;
        p.uid = st.inject_mode == PRB_INJECT_TRACE ? st.tr_uid[(int32_t) d.f_pos[f]] : d.f_pos[f];
        if (pi >= 0) post_send (st.L + src, pi, p);
        post_kind = 1; post_arg = f;
        return;
      }
    set_p_type (p, PRB_DATA);
    set_p_dst (p, dst); set_p_src (p, src); set_p_next_hop (p, src); set_p_last_hop (p, NODE_NONE);
    set_p_prev_node (p, NODE_NONE);
    set_p_szc (p, SZC_DATA_INJ);
    p.aux = (uint32_t) (uint64_t) ns_to_seconds (st, sc->now);
    int both = st.map_overlay[src] >= 0 && st.map_overlay[dst] >= 0;
    if (st.inject_mode == PRB_INJECT_TRACE)
      {
        int32_t i = (int32_t) d.f_pos[f];
        p.uid = st.tr_uid[i];
        set_p_valable (p, st.tr_flags[i] & 1);
        set_p_rejected (p, (st.tr_flags[i] >> 1) & 1);
      }
    else
      {
        p.uid = sc->next_uid++;
        set_p_valable (p, both ? 1 : 0);
        double prej = (st.opt_agent && st.opt_rejected) ? st.opt_rejected[src * st.N + dst] : 0.0;
        int rej = 0;
        if (prej > 0.0)
          {
            /* u in (0,1) < prej can only hold for prej > 0: the draw is skipped otherwise */
            uint64_t eid = st.env_id_base + (uint64_t) env;
            uint32_t r[4];
            philox4x32 ((uint32_t) eid, (uint32_t) (eid >> 32), (uint32_t) (src * st.N + dst), d.f_pos[f] - 1,
                        (uint32_t) st.seed, (uint32_t) (st.seed >> 32), r);
            double u = dmul_ (dadd_ ((double) r[2], 0.5), 1.0 / 4294967296.0);
            rej = u < prej ? 1 : 0;
          }
        set_p_rejected (p, rej);
      }
    set_p_first_hop (p, p_valable (p));
    set_p_ipdst (p, dst);
    if (pi >= 0) post_send (st.L + src, pi, p);
    post_kind = 1; post_arg = f;     /* ScheduleNextTx after the send (poisson-application.cc:357) */
  }

  /* ping timers are scheduled while the envs are built, Simulator::Stop right before Run
   * (sim.cc:545,703): they hold the lowest scheduler uids; then the sources start (sim.cc:599-632) */
  PRB_HD void
  start ()
  {
    sc->started = 1;
    for (int m = 0; m < st.M; m++) set_timer (slot_ping (st, m), st.ping_interval_ns);
    sc->stop_seq = sc->seq++;
    for (int f = 0; f < st.F; f++) schedule_flow (f, true);
  }

  /* executes the event in timer slot `slot` (the group's minimum), or resumes the suspended one */
  PRB_HD int
  handle (int slot, int64_t ts, uint32_t seq, int32_t action)
  {
    int status = ST_CONTINUE;
    ps_n = 0; ps_head = NIL; ps_tail = NIL; post_kind = 0;
    if (slot == SLOT_RESUME)
      resume (action);
    else
      {
        sc->now = ts;
        cur_seq = seq;
        sc->c.events++;
        if (ts - sc->epoch > (1ll << 31)) rebase (ts);
        if (slot < st.LA)
          transmit_complete (slot);
        else if (slot < 2 * st.LA)
          {
            int l = slot - st.LA;
            uint16_t h = d.w_head[l];
            uint16_t nx = d.pool[h].next;
            d.w_head[l] = nx;
            if (nx == NIL)
              {
                d.w_tail[l] = NIL;
                clear_timer (slot);
              }
            else
              put_timer (slot, d.pool[nx].rcv_ts, d.pool[nx].rcv_seq);
            status = receive (l, h);
          }
        else if (slot < 2 * st.LA + st.M)
          ping_timer (slot - 2 * st.LA);
        else
          inject (slot - 2 * st.LA - st.M);
      }
    flush_sends ();
    if (post_kind == 1)
      schedule_flow (post_arg, false);
    else if (post_kind == 2)
      set_timer (slot_ping (st, post_arg), sc->now + st.ping_interval_ns);
    return status;
  }
};

/* ---------------------------------------------------------------------------------------------
 * Next-event search.  Per event every lane of the group looks at its share of the 32-bit scan keys
 * (slots gl, gl+G, ...) and at its cached flow minimum; the group then reduces (key, uid).
 * ------------------------------------------------------------------------------------------- */

/* one lane's share of the key scan; (bk, bslot) start as (KEY_IDLE, -1) */
PRB_HD void
scan_keys (const Static &st, const Dyn &d, int gl, int G, uint32_t &bk, int &bslot)
{
  for (int s = gl; s < st.NKP; s += G)
    {
      uint32_t k = d.t_key[s];
      if (k < bk || (k == bk && k != KEY_IDLE && d.t_seq[s] < d.t_seq[bslot]))
        {
          bk = k;
          bslot = s;
        }
    }
}

/* Flow timers (exact 64-bit, one per source: up to N * (N - 1) of them) are kept under a two-level minimum: groups of 32
 * consecutive slots with their earliest timer in fg_ts / fg_seq / fg_slot, and every lane's share of the GROUP minima in
 * its registers.  A source that fires costs one rescan of its group (32 slots, shared by the lanes) and one pass of a lane
 * over its groups, instead of a pass of one lane over its share of ALL flows (310 slots at 100 nodes). */
PRB_HD void
fg_rescan_serial (const Static &st, const Dyn &d, int g)
{
  int64_t bt = T_IDLE; uint32_t bq = 0xFFFFFFFFu; int bs = -1;
  const int s0 = st.NK + 32 * g, s1 = s0 + 32 < st.S ? s0 + 32 : st.S;
  for (int s = s0; s < s1; s++)
    {
      int64_t t = d.t_ts[s];
      if (t == T_IDLE || t > bt) continue;
      uint32_t q = d.t_seq[s];
      if (t < bt || q < bq)
        {
          bt = t; bq = q; bs = s;
        }
    }
  d.fg_ts[g] = bt; d.fg_seq[g] = bq; d.fg_slot[g] = bs;
}

/* Few flows (Abilene: 110): a lane's share of ALL flow timers is a handful of slots and the plain scan is cheaper. */
PRB_HD bool
flows_two_level (const Static &st)
{
  return st.F > 256;
}

/* one lane's share of the flow timers (exact 64-bit): slots NK+gl, NK+gl+G, ... */
PRB_HD void
scan_flows (const Static &st, const Dyn &d, int gl, int G, int64_t &fts, uint32_t &fseq, int &fslot)
{
  fts = T_IDLE; fseq = 0xFFFFFFFFu; fslot = -1;
  for (int s = st.NK + gl; s < st.S; s += G)
    {
      int64_t t = d.t_ts[s];
      if (t == T_IDLE || t > fts) continue;
      uint32_t q = d.t_seq[s];
      if (t < fts || q < fseq)
        {
          fts = t; fseq = q; fslot = s;
        }
    }
}

/* one lane's share of the group minima: groups gl, gl+G, ... */
PRB_HD void
scan_flow_groups (const Static &st, const Dyn &d, int gl, int G, int64_t &fts, uint32_t &fseq, int &fslot)
{
  fts = T_IDLE; fseq = 0xFFFFFFFFu; fslot = -1;
  for (int g = gl; g < st.FG; g += G)
    {
      int64_t t = d.fg_ts[g];
      if (t == T_IDLE || t > fts) continue;
      uint32_t q = d.fg_seq[g];
      if (t < fts || q < fseq)
        {
          fts = t; fseq = q; fslot = d.fg_slot[g];
        }
    }
}

PRB_HD uint32_t
flow_key (const Scal *sc, int64_t fts)
{
  if (fts == T_IDLE) return KEY_IDLE;
  uint64_t dlt = (uint64_t) (fts - sc->epoch);
  return dlt < (uint64_t) KEY_FAR ? (uint32_t) dlt : KEY_FAR;
}

/* exact search over the 64-bit truth: only used when every armed timer is farther than the key range */
PRB_HD int
slow_min (const Static &st, const Dyn &d)
{
  int best = -1;
  int64_t bts = T_IDLE;
  uint32_t bseq = 0;
  for (int s = 0; s < st.S; s++)
    {
      if (s < st.NK && d.t_key[s] == KEY_IDLE) continue;
      int64_t t = d.t_ts[s];
      if (t == T_IDLE) continue;
      uint32_t q = d.t_seq[s];
      if (best < 0 || t < bts || (t == bts && q < bseq))
        {
          best = s; bts = t; bseq = q;
        }
    }
  return best;
}

/* TransmitComplete events that were elided and lie before the scheduler frontier: they count as executed */
PRB_HD int64_t
elided_before_frontier (const Static &st, const Dyn &d)
{
  const Scal *sc = d.sc;
  int64_t n = 0;
  for (int l = 0; l < st.LA; l++)
    {
      int64_t tc = d.t_ts[l];
      if (tc != T_IDLE && d.t_key[l] == KEY_IDLE && (tc < sc->front_ts || (tc == sc->front_ts && d.t_seq[l] < sc->front_seq)))
        n++;
    }
  return n;
}

/* ComputeStats::getAverage / getSignalingOverhead at read-out (compute-stats-v2.cc:87-97,263-266) */
PRB_HD void
finalize_counters (const Static &st, const Dyn &d, prb_counters &c)
{
  c = d.sc->c;
  c.events += elided_before_frontier (st, d);
  c.avg_e2e = c.n_e2e ? c.sum_e2e / (float) c.n_e2e : 0.0f;
  c.avg_cost = c.n_cost ? c.sum_cost / (float) c.n_cost : 0.0f;
  c.avg_underlay_e2e = c.n_underlay_e2e ? c.sum_underlay_e2e / (float) c.n_underlay_e2e : 0.0f;
  c.avg_underlay_cost = c.n_underlay_cost ? c.sum_underlay_cost / (float) c.n_underlay_cost : 0.0f;
  c.signaling_overhead = c.bytes_data ? (float) (int32_t) c.bytes_signaling / (float) (int32_t) c.bytes_data : 0.0f;
}

/* what the loop does once the next event (slot, ts, seq) is known; returns false when the env stops here */
template <bool LOG, bool GYM>
PRB_HD bool
step_or_stop (const Static &st, Sim<LOG, GYM> &sim, int slot, int64_t until_ns, int32_t action, int &status)
{
  Scal *sc = sim.sc;
  const int64_t stop = st.sim_time_ns;
  status = ST_CONTINUE;
  if (slot == SLOT_RESUME)
    {
      status = sim.handle (SLOT_RESUME, 0, 0, action);
      return status != ST_PAUSE;
    }
  if (slot < 0)
    {
      sc->finished = 1;
      sc->front_ts = stop; sc->front_seq = sc->stop_seq;
      return false;
    }
  int64_t ts = sim.d.t_ts[slot];
  uint32_t seq = sim.d.t_seq[slot];
  if (ts > stop || (ts == stop && seq > sc->stop_seq))
    {
      sc->now = stop;
      sc->finished = 1;
      sc->front_ts = stop; sc->front_seq = sc->stop_seq;
      return false;
    }
  if (ts >= until_ns && until_ns < stop)
    {
      if (until_ns > sc->front_ts)       /* the frontier never moves backwards (prb_run to an earlier time: a no-op) */
        {
          sc->front_ts = until_ns; sc->front_seq = 0;
        }
      return false;
    }
  status = sim.handle (slot, ts, seq, 0);
  return status != ST_PAUSE;
}

/* ---------------------------------------------------------------------------------------------
 * Serial driver (a group of one lane): the host test harness runs this; the kernels run the same
 * scan/handle functions with G lanes and warp reductions (prisma_b200.cu).
 * ------------------------------------------------------------------------------------------- */
template <bool LOG, bool GYM>
PRB_HDN inline void
run_env_serial (const Static &st, Dyn &d, int env, int64_t until_ns, int32_t action, prb_step_rec *out)
{
  Scal *sc = d.sc;
  prb_step_rec local;
  prb_step_rec *rec = (GYM && out) ? out : &local;
  Sim<LOG, GYM> sim (st, d, env, rec);
  if (GYM) rec->valid = 0;
  if (sc->finished) return;
  const bool fresh = !sc->started;
  if (fresh) sim.start ();
  bool resume_first = sc->pending != 0;
  const int64_t stop = st.sim_time_ns;
  int64_t fts; uint32_t fseq; int fslot;
  const bool two_level = flows_two_level (st);
  if (fresh && two_level)
    for (int g = 0; g < st.FG; g++) fg_rescan_serial (st, d, g);
  if (two_level)
    scan_flow_groups (st, d, 0, 1, fts, fseq, fslot);
  else
    scan_flows (st, d, 0, 1, fts, fseq, fslot);
  for (;;)
    {
      uint32_t bk = KEY_IDLE;
      int bslot = -1;
      if (resume_first)
        {
          int status;
          resume_first = false;
          if (!step_or_stop (st, sim, SLOT_RESUME, until_ns, action, status)) break;
          continue;
        }
      scan_keys (st, d, 0, 1, bk, bslot);
      uint32_t fk = flow_key (sc, fts);
      if (fk < bk || (fk == bk && fk != KEY_IDLE && fseq < d.t_seq[bslot]))
        {
          bk = fk;
          bslot = fslot;
        }
      if (bk == KEY_FAR)
        {
          /* everything armed is far away: move the epoch up to the clock, then fall back to the exact search */
          sim.rebase (sc->now);
          bslot = slow_min (st, d);
        }
      int status;
      bool fired_flow = bslot >= st.NK;
      if (!step_or_stop (st, sim, bslot, until_ns, 0, status)) break;
      if (fired_flow && two_level)
        {
          fg_rescan_serial (st, d, (bslot - st.NK) >> 5);
          scan_flow_groups (st, d, 0, 1, fts, fseq, fslot);
        }
      else if (fired_flow)
        scan_flows (st, d, 0, 1, fts, fseq, fslot);
    }
  if (!sc->pending && until_ns < stop && sc->now < until_ns && !sc->finished) sc->now = until_ns;
  sc->c.end_ns = sc->now;
}

}  // namespace prb
