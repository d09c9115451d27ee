/* TEST INFRASTRUCTURE — see prisma_oracle.h.  CPU restatement of the PRISMA/ns-3 step loop.
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared (oracle/Makefile).  x86-64 only
 * (ns-3's int64x64_t(long double) constructor relies on the 80-bit x87 format).
 */
#include "prisma_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------------------------------
 * ns-3 time arithmetic
 * ---------------------------------------------------------------------------------------- */

/* Seconds(double).GetNanoSeconds():
 *   int64x64_t(long double) — ns3-gym/src/core/model/int64x64-128.h:94-126 (split with modf,
 *   fraction scaled by 2^64 and rounded with +0.5, carry on roll-over);
 *   Time::From multiplies by the integer factor 1e9 — nstime.h:456-475, int64x64-128.cc:64-110;
 *   Time(int64x64_t) keeps the high word — nstime.h:235-236. */
int64_t
po_seconds_to_ns (double s)
{
  int neg = s < 0;
  long double v = neg ? -(long double) s : (long double) s;
  long double fhi;
  long double flo = modfl (v, &fhi);
  const long double two64 = 18446744073709551616.0L;
  flo = flo * two64 + 0.5L;
  u128 hi = (u128) fhi;
  uint64_t lo = (uint64_t) flo;
  if (flo >= two64)
    hi += 1;
  /* Q64.64 times the integer 1e9: hi*1e9 + (lo*1e9 >> 64), fraction dropped by GetHigh */
  u128 mid = (u128) lo * 1000000000u;
  u128 q_hi = hi * 1000000000u + (mid >> 64);
  uint64_t q_lo = (uint64_t) mid;
  if (!neg)
    return (int64_t) q_hi;
  /* negate the 128-bit value, then arithmetic shift (floor) */
  if (q_lo != 0)
    return -(int64_t) q_hi - 1;
  return -(int64_t) q_hi;
}

/* int64x64_t::Udiv (1.0, v) as used by int64x64_t::Invert — int64x64-128.cc:121-186,216-232.
 * Long division that produces 64 fraction bits in rounds, shedding denominator bits when it
 * has to; restated here for a = 2^64 (the Q64.64 one) and the integer v. */
static u128
po_udiv_q (u128 a, u128 b)
{
  u128 rem = a, den = b;
  u128 quo = rem / den;
  rem = rem % den;
  u128 result = quo;
  const unsigned DIGITS = 64;
  const u128 HIBIT = (u128) 1 << 127;
  unsigned digis = 0, shift = 0;
  while (shift < DIGITS && !(den & 1))
    {
      shift++;
      den >>= 1;
    }
  while (digis < DIGITS && rem != 0)
    {
      while (digis + shift < DIGITS && !(rem & HIBIT))
        {
          shift++;
          rem <<= 1;
        }
      while (digis + shift < DIGITS && (!(den & 1) || rem < den))
        {
          shift++;
          den >>= 1;
        }
      quo = rem / den;
      rem = rem % den;
      result <<= shift;
      result += quo;
      digis += shift;
      shift = 0;
    }
  if (digis < DIGITS)
    result <<= (DIGITS - digis);
  return result;
}

static u128
po_umul_by_invert (u128 a, u128 b)
{
  /* int64x64-128.cc:200-214 */
  u128 ah = a >> 64, bh = b >> 64;
  u128 al = a & 0xFFFFFFFFFFFFFFFFull, bl = b & 0xFFFFFFFFFFFFFFFFull;
  u128 hi = ah * bh;
  u128 mid = ah * bl + al * bh;
  mid >>= 64;
  return hi + mid;
}

static u128
po_invert (uint64_t v)
{
  /* int64x64_t::Invert — int64x64-128.cc:216-232 */
  u128 one = (u128) 1 << 64;
  u128 r = po_udiv_q (one, (u128) v);
  u128 tmp = po_umul_by_invert ((u128) v << 64, r);
  if ((uint64_t) (tmp >> 64) != 1)
    r += 1;
  return r;
}

void
po_invert_1e9 (uint64_t *hi, uint64_t *lo)
{
  u128 r = po_invert (1000000000ull);
  *hi = (uint64_t) (r >> 64);
  *lo = (uint64_t) r;
}

/* Time::GetSeconds(): To(S) = MulByInvert(Invert(1e9)) then int64x64_t::GetDouble
 * — nstime.h:505-523, int64x64-128.cc:191-214, int64x64-128.h:203-213 (long double sum, then
 * narrowed to double: two roundings). */
double
po_ns_to_seconds (int64_t ns)
{
  static u128 inv = 0;
  if (inv == 0)
    inv = po_invert (1000000000ull);
  int neg = ns < 0;
  u128 a = (u128) (uint64_t) (neg ? -ns : ns) << 64;
  u128 q = po_umul_by_invert (a, inv);
  long double fhi = (long double) (uint64_t) (q >> 64);
  long double flo = (long double) (uint64_t) q / 18446744073709551616.0L;
  long double r = fhi + flo;
  if (neg)
    r = -r;
  return (double) r;
}

/* DataRate::CalculateBytesTxTime — ns3-gym/src/network/utils/data-rate.cc:235-240 */
int64_t
po_tx_time_ns (uint32_t bytes, uint64_t bps)
{
  double s = (double) bytes * 8 / (double) bps;
  return po_seconds_to_ns (s);
}

/* ------------------------------------------------------------------------------------------
 * Philox-4x32-10
 * ---------------------------------------------------------------------------------------- */
void
po_philox4x32 (const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
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

/* -log(u) for u in (0,1), built from IEEE-754 +,-,*,/ and fma only so that the CUDA path
 * (which uses the _rn intrinsics in the same order) reproduces it bit for bit.  This is the
 * synthetic-injection spec of prisma_b200 (DESIGN.md "Synthetic traffic"), not ns-3's libm call
 * (random-variable-stream.cc:406-428 uses std::log on an MRG32k3a draw; parity with that is by
 * trace replay). */
static double
po_neglog (double u)
{
  uint64_t bits;
  memcpy (&bits, &u, 8);
  int e = (int) ((bits >> 52) & 0x7FF) - 1023;
  uint64_t mb = (bits & 0xFFFFFFFFFFFFFull) | 0x3FF0000000000000ull;
  double m;
  memcpy (&m, &mb, 8);                  /* m in [1,2) */
  if (m > 1.4142135623730951)
    {
      m = m * 0.5;
      e += 1;
    }
  double s = (m - 1.0) / (m + 1.0);
  double z = s * s;
  /* 2*atanh(s) = 2s(1 + z/3 + z^2/5 + ...), |s| <= 0.1716 */
  double p = 1.0 / 27.0;
  p = fma (p, z, 1.0 / 25.0);
  p = fma (p, z, 1.0 / 23.0);
  p = fma (p, z, 1.0 / 21.0);
  p = fma (p, z, 1.0 / 19.0);
  p = fma (p, z, 1.0 / 17.0);
  p = fma (p, z, 1.0 / 15.0);
  p = fma (p, z, 1.0 / 13.0);
  p = fma (p, z, 1.0 / 11.0);
  p = fma (p, z, 1.0 / 9.0);
  p = fma (p, z, 1.0 / 7.0);
  p = fma (p, z, 1.0 / 5.0);
  p = fma (p, z, 1.0 / 3.0);
  p = fma (p, z, 1.0);
  double lm = 2.0 * s * p;
  double r = fma ((double) e, 0.6931471805599453, lm);
  return -r;
}

static double
po_u01_53 (uint32_t a, uint32_t b)
{
  /* (0,1): 52-bit integer + 0.5 (exact in binary64), scaled by 2^-52 */
  uint64_t x = (((uint64_t) a << 32) | b) >> 12;
  return ((double) x + 0.5) * (1.0 / 4503599627370496.0);
}

/* ------------------------------------------------------------------------------------------
 * Topology and routing
 * ---------------------------------------------------------------------------------------- */

int32_t
po_num_links (int32_t n, const uint8_t *adj)
{
  int32_t e = 0;
  for (int i = 0; i < n; i++)
    for (int j = i + 1; j < n; j++)
      if (adj[i * n + j])
        e++;
  return 2 * e;
}

/* sim.cc:414-433: one p2p link per 1 in the upper triangle, device on i created first, then on j. */
int32_t
po_build_links (int32_t n, const uint8_t *adj, int32_t *from, int32_t *to)
{
  int32_t l = 0;
  for (int i = 0; i < n; i++)
    for (int j = i + 1; j < n; j++)
      if (adj[i * n + j])
        {
          from[l] = i; to[l] = j; l++;
          from[l] = j; to[l] = i; l++;
        }
  return l;
}

/* Next hop of ns-3 global routing as PRISMA uses it.
 * GlobalRouteManagerImpl::SPFCalculate/SPFNext (global-route-manager-impl.cc:747-980) runs a
 * unit-metric SPF per root; equal-cost parents merge their root-exit lists, which are kept sorted by
 * (next-hop address, interface) (SPFVertex::MergeRootExitDirections, :321-333); one host route per
 * exit is installed in that order (SPFIntraAddRouter, :1342-1440) and Ipv4GlobalRouting::LookupGlobal
 * returns the first match when random ECMP is off (ipv4-global-routing.cc:139-235).  Next-hop
 * addresses are handed out in device-creation order (sim.cc:442-447), so for a root the order of its
 * neighbours' addresses is the order of edge creation, i.e. ascending neighbour id.  Result: among
 * all shortest paths the first hop with the smallest node id wins. */
int32_t
po_global_routes (int32_t n, const uint8_t *adj, int32_t *route_out)
{
  int32_t nl = po_num_links (n, adj);
  int32_t *from = (int32_t *) malloc (sizeof (int32_t) * (nl + 1));
  int32_t *to = (int32_t *) malloc (sizeof (int32_t) * (nl + 1));
  po_build_links (n, adj, from, to);
  int32_t *dist = (int32_t *) malloc (sizeof (int32_t) * n);
  int32_t *queue = (int32_t *) malloc (sizeof (int32_t) * n);
  for (int d = 0; d < n; d++)
    {
      /* BFS from the destination (links are symmetric) */
      for (int i = 0; i < n; i++) dist[i] = -1;
      int qh = 0, qt = 0;
      dist[d] = 0; queue[qt++] = d;
      while (qh < qt)
        {
          int v = queue[qh++];
          for (int w = 0; w < n; w++)
            if ((adj[v * n + w] || adj[w * n + v]) && v != w && dist[w] < 0)
              {
                dist[w] = dist[v] + 1;
                queue[qt++] = w;
              }
        }
      for (int x = 0; x < n; x++)
        {
          route_out[x * n + d] = -1;
          if (x == d || dist[x] < 0) continue;
          for (int y = 0; y < n && route_out[x * n + d] < 0; y++)
            {
              if (y == x || !(adj[x * n + y] || adj[y * n + x])) continue;
              if (dist[y] != dist[x] - 1) continue;
              for (int l = 0; l < nl; l++)
                if (from[l] == x && to[l] == y)
                  {
                    route_out[x * n + d] = l;
                    break;
                  }
            }
        }
    }
  free (from); free (to); free (dist); free (queue);
  return nl;
}

/* ------------------------------------------------------------------------------------------
 * Simulator state
 * ---------------------------------------------------------------------------------------- */

typedef struct po_pkt
{
  uint8_t type, valable, rejected, first_hop; /* first_hop: overlay packet still addressed to its own switch (sim.cc:615) */
  int32_t src, dst, last_hop, next_hop;
  int32_t ipdst;          /* node d whose traffic-node address 10.2.2.(d+1) the IP header carries */
  uint32_t size;          /* wire bytes incl. the 2-byte PPP header */
  int64_t uid;
  uint64_t start;         /* tag start time: whole seconds (data) or milliseconds (ping forward) */
  int64_t birth_ns;
  float one_hop_delay;    /* ping back */
  uint32_t ping_idx, tunnel_idx;
  int64_t pkt_id;         /* small signalling: uid of the data packet it reports */
  uint32_t seg_index, nn_index; /* big signalling: m_segIndex / m_NNIndex (big-signaling-application.cc:241-246,276-277) */
  /* restated Python bookkeeping: Agent.temp_obs / pkt_tracking_dict (forwarder.py:153-159,334-350) */
  int32_t tracked;
  int32_t prev_node;
  int64_t prev_time_ns;
  struct po_pkt *next;    /* FIFO link inside a queue */
} po_pkt;

enum { EVK_INJ = 0, EVK_TXC = 1, EVK_RCV = 2, EVK_PING = 3 };

typedef struct
{
  int64_t ts;
  uint64_t seq;
  int32_t kind;
  int32_t a;      /* flow / link / overlay node */
  po_pkt *pkt;
} po_hev;

typedef struct
{
  int32_t from, to;       /* node ids; access link: from = N+i (traffic node), to = i */
  uint64_t bps;
  int64_t delay_ns;
  int32_t byte_mode;      /* 1: MaxSize in bytes (sim.cc:424); 0: 1000 packets (sim.cc:405) */
  uint32_t max_size;
  uint32_t q_bytes, q_pkts;
  po_pkt *q_head, *q_tail;
  int32_t busy;
} po_link;

typedef struct
{
  int32_t src, dst;
  uint64_t bps;
  double mean_s;
  uint64_t n_draw;        /* Philox draw ordinal */
  int32_t big;            /* 1: BigSignalingGeneratorApplication (fixed gap), 0: Poisson source */
  uint32_t seg_index, nn_index;
  int64_t trace_pos;      /* replay: index of this flow's next injection in the trace */
} po_flow;

typedef struct
{
  uint32_t idx;
  uint64_t start_ms;
} po_sent_ping;

typedef struct
{
  int32_t node;           /* underlay id */
  int32_t k;
  int32_t nbr[PO_MAX_K];  /* underlay ids of overlay neighbours, action order (sim.cc:469-476) */
  uint32_t ping_index;    /* m_pingPacketIndex */
  po_sent_ping *sent[PO_MAX_K]; int32_t n_sent[PO_MAX_K], cap_sent[PO_MAX_K];
  float *win[PO_MAX_K]; int32_t n_win[PO_MAX_K];
  int64_t *lost; int32_t n_lost, cap_lost;
  uint32_t sign_size;     /* smallSignalingSize (sim.cc:372-393) */
  int32_t last_data_dst;  /* DataPacketManager::m_destination: survives until the next data packet */
} po_agent;

struct po_sim
{
  po_config cfg;
  int32_t N, M, L;        /* L switch<->switch directed links; access links are L..L+N-1 */
  po_link *links;
  int32_t *route;         /* [N*N] */
  int32_t *o2u;           /* [M] */
  int32_t *map_overlay;   /* [N] */
  int32_t *degree;        /* [N] */
  po_agent *agents;
  po_flow *flows; int32_t F;
  int32_t *flow_of;       /* [N*N] flow index or -1 */
  int32_t *big_flow_of;   /* [N*N] big-signalling source index or -1 */
  int64_t big_start_ns, big_gap_ns; uint32_t big_segs;
  double *opt_rejected;
  /* scheduler */
  po_hev *heap; int64_t heap_n, heap_cap;
  uint64_t seq, stop_seq;
  int64_t now;
  /* injection trace */
  int64_t tr_n; int64_t *tr_t; int32_t *tr_src, *tr_dst; int64_t *tr_uid; uint8_t *tr_val, *tr_rej; int64_t *tr_next;
  int64_t next_uid;
  /* policy */
  int32_t policy; int32_t *pol_data; int64_t pol_n, pol_pos; uint64_t pol_rng;
  /* logs */
  int32_t log_events;
  po_event *ev; int64_t ev_n, ev_cap;
  po_decision *dec; int64_t dec_n, dec_cap;
  po_counters c;
  int32_t started;
};

static void
heap_push (po_sim *s, int64_t ts, int32_t kind, int32_t a, po_pkt *pkt)
{
  if (s->heap_n == s->heap_cap)
    {
      s->heap_cap = s->heap_cap ? s->heap_cap * 2 : 1024;
      s->heap = (po_hev *) realloc (s->heap, sizeof (po_hev) * s->heap_cap);
    }
  po_hev e = { ts, s->seq++, kind, a, pkt };
  int64_t i = s->heap_n++;
  while (i > 0)
    {
      int64_t p = (i - 1) / 2;
      po_hev *q = &s->heap[p];
      /* ns-3 event key: (timestamp, uid) — core/model/scheduler.h:153-169 */
      if (q->ts < e.ts || (q->ts == e.ts && q->seq < e.seq)) break;
      s->heap[i] = *q;
      i = p;
    }
  s->heap[i] = e;
}

static po_hev
heap_pop (po_sim *s)
{
  po_hev top = s->heap[0];
  po_hev e = s->heap[--s->heap_n];
  int64_t i = 0;
  for (;;)
    {
      int64_t c = 2 * i + 1;
      if (c >= s->heap_n) break;
      if (c + 1 < s->heap_n)
        {
          po_hev *a = &s->heap[c], *b = &s->heap[c + 1];
          if (b->ts < a->ts || (b->ts == a->ts && b->seq < a->seq)) c++;
        }
      po_hev *q = &s->heap[c];
      if (e.ts < q->ts || (e.ts == q->ts && e.seq < q->seq)) break;
      s->heap[i] = *q;
      i = c;
    }
  if (s->heap_n > 0) s->heap[i] = e;
  return top;
}

static void
log_event (po_sim *s, int kind, int node, int link, const po_pkt *p, int qbytes)
{
  if (!s->log_events) return;
  if (s->ev_n == s->ev_cap)
    {
      s->ev_cap = s->ev_cap ? s->ev_cap * 2 : 4096;
      s->ev = (po_event *) realloc (s->ev, sizeof (po_event) * s->ev_cap);
    }
  po_event *e = &s->ev[s->ev_n++];
  e->t_ns = s->now; e->kind = kind; e->node = node; e->link = link;
  e->uid = p->type == PO_DATA ? p->uid : -1;
  e->size = (int32_t) p->size; e->type = p->type; e->src = p->src; e->dst = p->dst;
  e->last_hop = p->last_hop; e->next_hop = p->next_hop; e->valable = p->type == PO_DATA ? p->valable : -1;
  e->qbytes = qbytes;
}

/* ------------------------------------------------------------------------------------------
 * Devices: Send / TransmitStart / TransmitComplete
 * ---------------------------------------------------------------------------------------- */

static po_pkt *
pkt_clone (const po_pkt *p)
{
  po_pkt *q = (po_pkt *) malloc (sizeof (po_pkt));
  *q = *p;
  q->next = 0;
  return q;
}

static void
push_f (float *sum, int64_t *n, float v)
{
  /* ComputeStats keeps float vectors and re-sums them front to back in float
   * (compute-stats-v2.cc:87-97); a running float sum in push order is the same value. */
  *sum += v;
  *n += 1;
}

/* MacTxDrop fan-out: every agent's DataPacketManager::dropPacket is connected to every
 * switch<->switch device (data-packet-manager.cc:100-110); the one whose node is the tag's
 * lastHop keeps the uid (data-packet-manager.cc:87-98). */
static void
mac_tx_drop (po_sim *s, int32_t link, const po_pkt *p)
{
  if (link >= s->L) return;                 /* devices 0/1 are not hooked */
  if (p->type != PO_DATA || !p->valable) return;
  for (int m = 0; m < s->M; m++)
    {
      po_agent *a = &s->agents[m];
      if (p->dst != a->node && p->last_hop == a->node)
        {
          if (a->n_lost == a->cap_lost)
            {
              a->cap_lost = a->cap_lost ? a->cap_lost * 2 : 16;
              a->lost = (int64_t *) realloc (a->lost, sizeof (int64_t) * a->cap_lost);
            }
          a->lost[a->n_lost++] = p->uid;
        }
    }
}

static void transmit_start (po_sim *s, int32_t l, po_pkt *p);

/* PointToPointNetDevice::Send — prisma/ns3/point-to-point-net-device.cc:594-666;
 * Queue<Item>::DoEnqueue — ns3-gym/src/network/utils/queue.h:444-468.  Takes ownership of p. */
static void
dev_send (po_sim *s, int32_t l, po_pkt *p)
{
  po_link *k = &s->links[l];
  int32_t node = k->from;
  if (p->type == PO_DATA && p->rejected == 1 && node < 23)
    {
      if (p->valable)
        {
          s->c.overlay_lost++;
          push_f (&s->c.sum_cost, &s->c.n_cost, (float) s->cfg.loss_penalty);
        }
      else
        {
          push_f (&s->c.sum_underlay_cost, &s->c.n_underlay_cost, (float) s->cfg.loss_penalty);
          s->c.underlay_lost++;
        }
      mac_tx_drop (s, l, p);
      free (p);
      return;
    }
  if (l >= s->L) log_event (s, PO_EV_INJ, node, l, p, -1);   /* MacTx on the traffic node's device */
  int fits = k->byte_mode ? (k->q_bytes + p->size <= k->max_size) : (k->q_pkts + 1 <= k->max_size);
  if (!fits)
    {
      if (l < s->L) log_event (s, PO_EV_DRP, node, l, p, (int) k->q_bytes);
      if (p->type == PO_DATA)
        {
          if (p->valable && p->dst != node)
            {
              s->c.overlay_lost++;
              push_f (&s->c.sum_cost, &s->c.n_cost, (float) s->cfg.loss_penalty);
            }
          else
            {
              push_f (&s->c.sum_underlay_cost, &s->c.n_underlay_cost, (float) s->cfg.loss_penalty);
              s->c.underlay_lost++;
            }
          mac_tx_drop (s, l, p);
        }
      free (p);
      return;
    }
  p->next = 0;
  if (k->q_tail) k->q_tail->next = p; else k->q_head = p;
  k->q_tail = p;
  k->q_bytes += p->size;
  k->q_pkts += 1;
  if (l < s->L) log_event (s, PO_EV_ENQ, node, l, p, (int) k->q_bytes);
  if (!k->busy)
    {
      po_pkt *h = k->q_head;
      k->q_head = h->next;
      if (!k->q_head) k->q_tail = 0;
      k->q_bytes -= h->size;
      k->q_pkts -= 1;
      transmit_start (s, l, h);
    }
}

/* PointToPointNetDevice::TransmitStart — point-to-point-net-device.cc:272-302 (TransmitComplete is
 * scheduled before the channel schedules the peer's Receive);
 * PointToPointChannel::TransmitStart — ns3-gym/src/point-to-point/model/point-to-point-channel.cc:86-107 */
static void
transmit_start (po_sim *s, int32_t l, po_pkt *p)
{
  po_link *k = &s->links[l];
  k->busy = 1;
  if (l < s->L) log_event (s, PO_EV_TXB, k->from, l, p, (int) k->q_bytes);
  int64_t tx = po_tx_time_ns (p->size, k->bps);
  heap_push (s, s->now + tx, EVK_TXC, l, 0);
  heap_push (s, s->now + tx + k->delay_ns, EVK_RCV, l, p);
}

/* PointToPointNetDevice::TransmitComplete — point-to-point-net-device.cc:304-336 */
static void
transmit_complete (po_sim *s, int32_t l)
{
  po_link *k = &s->links[l];
  k->busy = 0;
  po_pkt *h = k->q_head;
  if (!h) return;
  k->q_head = h->next;
  if (!k->q_head) k->q_tail = 0;
  k->q_bytes -= h->size;
  k->q_pkts -= 1;
  transmit_start (s, l, h);
}

/* ------------------------------------------------------------------------------------------
 * Agents: observation, Gym step, apply action, pings, signalling
 * ---------------------------------------------------------------------------------------- */

static uint64_t
now_ms (po_sim *s)
{
  /* Time::GetMilliSeconds — nstime.h ToInteger: integer division of the ns count */
  return (uint64_t) (s->now / 1000000);
}

/* PingBackPacketManager::getMaxTimePingForwardPacketSent — ping-back-packet-manager.cc:111-118 */
static float
ping_oldest_age (po_sim *s, po_agent *a, int t)
{
  if (a->n_sent[t] > 0)
    {
      double v = po_ns_to_seconds (s->now) - (double) a->sent[t][0].start_ms * 0.001;
      return (float) (v < 2.60 ? v : 2.60);
    }
  return 0.0f;
}

/* DataPacketManager::getObservation — data-packet-manager.cc:170-206 */
static int
build_obs (po_sim *s, po_agent *a, const po_pkt *p, int32_t *obs)
{
  obs[0] = s->map_overlay[p->dst];
  for (int i = 0; i < a->k; i++)
    {
      int32_t out = s->route[a->node * s->N + a->nbr[i]];
      uint32_t value;
      if (!s->cfg.ping_as_obs)
        value = s->links[out].q_bytes;
      else
        {
          double avg = 0.0;
          if (a->n_win[i] > 0)
            {
              double sum = 0.0;
              for (int j = 0; j < a->n_win[i]; j++) sum += (double) a->win[i][j];
              avg = sum / a->n_win[i];
            }
          double age = (double) ping_oldest_age (s, a, i);
          value = (uint32_t) (1000 * (avg > age ? avg : age));
        }
      obs[1 + i] = (int32_t) value;
    }
  return 1 + a->k;
}

static po_decision *
new_decision (po_sim *s)
{
  if (s->dec_n == s->dec_cap)
    {
      s->dec_cap = s->dec_cap ? s->dec_cap * 2 : 4096;
      s->dec = (po_decision *) realloc (s->dec, sizeof (po_decision) * s->dec_cap);
    }
  po_decision *d = &s->dec[s->dec_n++];
  memset (d, 0, sizeof (*d));
  return d;
}

static int32_t
policy_action (po_sim *s, int m, const int32_t *obs, int k)
{
  switch (s->policy)
    {
    case PO_POLICY_TABLE:   /* e.g. the sp agent: neighbors.index(shortest_path(G, me, dst)[1]) forwarder.py:190-191 */
      return s->pol_data[m * s->M + obs[0]];
    case PO_POLICY_TRACE:
      if (s->pol_pos >= s->pol_n) { fprintf (stderr, "prisma_oracle: action trace exhausted\n"); abort (); }
      return s->pol_data[s->pol_pos++];
    case PO_POLICY_RANDOM:  /* same xorshift64* as oracle/ref_harness/ref_harness.cc NextRand */
      s->pol_rng ^= s->pol_rng >> 12; s->pol_rng ^= s->pol_rng << 25; s->pol_rng ^= s->pol_rng >> 27;
      return (int32_t) ((s->pol_rng * 2685821657736338717ULL) % (uint64_t) k);
    default:
      return 0;
    }
}

/* DataPacketManager::sendSmallSignalingPacket — data-packet-manager.cc:301-347 */
static void
send_small_signaling (po_sim *s, po_agent *a, const po_pkt *p, int32_t in_link)
{
  if (p->src == a->node) return;
  po_pkt *q = (po_pkt *) calloc (1, sizeof (po_pkt));
  q->type = PO_SMALL_SIGN;
  q->dst = p->last_hop; q->next_hop = p->last_hop; q->last_hop = a->node; q->src = a->node;
  q->ipdst = p->last_hop;
  q->pkt_id = p->uid; q->uid = -1;
  q->size = a->sign_size + 8 + 20 + 2;
  q->birth_ns = s->now;
  dev_send (s, in_link ^ 1, q);   /* m_receivingNetDev->Send: the reverse direction of the arrival link */
}

/* One Gym step for a data packet: OpenGymInterface::NotifyCurrentState order
 * (ns3-gym/src/opengym/model/opengym_interface.cc:213-286) + Forwarder.step/run bookkeeping
 * (prisma/source/forwarder.py:135-161,291-379) + PacketRoutingEnv::ExecuteActions
 * (packet-routing-gym.cc:196-213) + DataPacketManager::sendPacket (data-packet-manager.cc:250-299). */
static void
agent_data_step (po_sim *s, int m, po_pkt *p, int32_t in_link)
{
  po_agent *a = &s->agents[m];
  po_decision *d = new_decision (s);
  d->t_ns = s->now; d->node = a->node; d->overlay = m; d->uid = p->uid; d->ptype = PO_DATA;
  d->nobs = build_obs (s, a, p, d->obs);
  d->done = p->dst == a->node;
  d->e2e_delay = (float) (po_ns_to_seconds (s->now) - (double) p->start);
  /* DataPacketManager::getInfo flushes the lost list back to front — data-packet-manager.cc:229-237 */
  d->n_lost = a->n_lost;
  for (int i = 0; i < a->n_lost && i < 8; i++) d->lost_uid[i] = a->lost[a->n_lost - 1 - i];
  a->n_lost = 0;
  s->c.env_steps++;
  /* Forwarder bookkeeping */
  d->is_new = !p->tracked;
  d->prev_node = -1;
  if (p->tracked)
    {
      d->prev_node = p->prev_node;
      d->hop_time = (double) (s->now - p->prev_time_ns) * 1e-9;
    }
  int consult = !d->done;   /* forwarder.py:149: obs[0]==self.index -> action 0, not recorded */
  int32_t action = 0;
  if (consult)
    {
      action = policy_action (s, m, d->obs, a->k);
      s->c.decisions++;
    }
  d->action = consult ? action : -1;
  /* ExecuteActions */
  if (s->cfg.train) send_small_signaling (s, a, p, in_link);
  if (d->done)
    return;
  if (action >= 0 && action < a->k)
    {
      po_pkt *q = pkt_clone (p);
      q->last_hop = a->node;
      q->next_hop = a->nbr[action];
      q->ipdst = a->nbr[action];
      q->first_hop = 0;
      q->size = s->cfg.packet_size + 8 + 20 + 2;
      q->tracked = 1; q->prev_node = a->node; q->prev_time_ns = s->now;
      dev_send (s, s->route[a->node * s->N + a->nbr[action]], q);
    }
  else
    s->c.vanished++;   /* data-packet-manager.cc:294-297: the packet silently disappears */
}

static void
agent_control_step (po_sim *s, int m, po_pkt *p)
{
  /* obs [1000], no action effect — packet-routing-gym.cc:150-162,196-213 */
  po_agent *a = &s->agents[m];
  po_decision *d = new_decision (s);
  d->t_ns = s->now; d->node = a->node; d->overlay = m; d->uid = p->type == PO_SMALL_SIGN ? p->pkt_id : -1;
  d->ptype = p->type; d->nobs = 1; d->obs[0] = 1000; d->action = -1; d->prev_node = -1;
  if (p->type == PO_BIG_SIGN)
    {
      /* info fields 18-20: NN Index, segment Index, NodeId Signaled (big-signaling-packet-manager.cc:104-121) */
      d->uid = (int64_t) (((uint64_t) p->nn_index << 32) | p->seg_index);
      d->prev_node = p->src;
    }
  /* GetGameOver asks the DATA manager, whose m_destination still holds the last data packet seen at
   * this node (packet-routing-gym.cc:142-147, data-packet-manager.cc:224-227) */
  d->done = a->last_data_dst == a->node;
  s->c.env_steps++;
}

/* PingForwardPacketManager::receivePacket/sendPingBackPacket — ping-forward-packet-manager.cc:93-156 */
static void
ping_forward_rx (po_sim *s, po_agent *a, const po_pkt *p, int32_t in_link)
{
  float delay = (float) (po_ns_to_seconds (s->now) - ((double) p->start * 0.001));
  po_pkt *q = (po_pkt *) calloc (1, sizeof (po_pkt));
  q->type = PO_PING_BACK;
  q->dst = p->last_hop; q->next_hop = p->last_hop; q->last_hop = a->node; q->src = a->node;
  q->ipdst = p->last_hop;
  q->one_hop_delay = delay; q->ping_idx = p->ping_idx; q->tunnel_idx = p->tunnel_idx;
  q->uid = -1; q->size = 8 + 8 + 20 + 2; q->birth_ns = s->now;
  dev_send (s, in_link ^ 1, q);
}

/* PingBackPacketManager::receivePacket — ping-back-packet-manager.cc:122-145 */
static void
ping_back_rx (po_sim *s, po_agent *a, const po_pkt *p)
{
  int t = (int) p->tunnel_idx;
  if (t >= PO_MAX_K) return;
  for (int i = 0; i < a->n_sent[t]; i++)
    if (a->sent[t][i].idx == p->ping_idx)
      {
        memmove (&a->sent[t][i], &a->sent[t][i + 1], sizeof (po_sent_ping) * (a->n_sent[t] - i - 1));
        a->n_sent[t]--;
        break;
      }
  uint32_t ma = s->cfg.moving_average_size;
  if ((uint32_t) a->n_win[t] >= ma && a->n_win[t] > 0)
    {
      memmove (&a->win[t][0], &a->win[t][1], sizeof (float) * (a->n_win[t] - 1));
      a->n_win[t]--;
    }
  a->win[t] = (float *) realloc (a->win[t], sizeof (float) * (a->n_win[t] + 1));
  a->win[t][a->n_win[t]++] = p->one_hop_delay;
}

/* DataPacketManager::sendPingPackets/sendPingForwardPacket — data-packet-manager.cc:349-413;
 * PingBackPacketManager::addSentPingForwardPacket — ping-back-packet-manager.cc:98-109 */
static void
ping_timer (po_sim *s, int m)
{
  po_agent *a = &s->agents[m];
  uint64_t ms = now_ms (s);
  for (int t = 0; t < a->k; t++)
    {
      if (a->n_sent[t] == a->cap_sent[t])
        {
          a->cap_sent[t] = a->cap_sent[t] ? a->cap_sent[t] * 2 : 16;
          a->sent[t] = (po_sent_ping *) realloc (a->sent[t], sizeof (po_sent_ping) * a->cap_sent[t]);
        }
      a->sent[t][a->n_sent[t]].idx = a->ping_index;
      a->sent[t][a->n_sent[t]].start_ms = ms;
      a->n_sent[t]++;
    }
  for (int t = 0; t < a->k; t++)
    {
      po_pkt *q = (po_pkt *) calloc (1, sizeof (po_pkt));
      q->type = PO_PING_FWD;
      q->dst = a->nbr[t]; q->next_hop = a->nbr[t]; q->last_hop = a->node; q->src = a->node;
      q->ipdst = a->nbr[t];
      q->start = ms; q->tunnel_idx = (uint32_t) t; q->ping_idx = a->ping_index;
      q->uid = -1; q->size = 8 + 8 + 20 + 2; q->birth_ns = s->now;
      dev_send (s, s->route[a->node * s->N + a->nbr[t]], q);
    }
  a->ping_index += 1;
  heap_push (s, s->now + s->cfg.ping_interval_ns, EVK_PING, m, 0);
}

/* ------------------------------------------------------------------------------------------
 * Receive, injection, event loop
 * ---------------------------------------------------------------------------------------- */

/* PointToPointNetDevice::Receive — point-to-point-net-device.cc:370-467, with the MacRx hook
 * PacketRoutingEnv::NotifyPktRcv — packet-routing-gym.cc:230-267 and the IP forwarding that
 * m_rxCallback leads to (patched Ipv4Interface::Send, prisma/ns3_model/ipv4-interface.cc:213-232). */
static void
dev_receive (po_sim *s, int32_t l, po_pkt *p)
{
  po_link *k = &s->links[l];
  int32_t x = k->to;
  int access = l >= s->L;
  if (p->type == PO_DATA && p->rejected == 1 && x < 23)
    {
      mac_tx_drop (s, access ? l : (l ^ 1), p);   /* fires on the receiving device */
      free (p);
      return;
    }
  log_event (s, access ? PO_EV_ARR : PO_EV_RCV, x, l, p, -1);
  if (!access)
    {
      s->c.packet_hops++;
      if (p->type == PO_DATA) s->c.data_hops++;
    }
  /* --- MacRx trace: agent hook (every device of an overlay node, sim.cc:554-557) --- */
  int m = s->map_overlay[x];
  if (m >= 0)
    {
      po_agent *a = &s->agents[m];
      switch (p->type)
        {
        case PO_DATA:
          /* PacketManager::receivePacket valid test (packet-manager.cc:116) and trafficValable
           * (data-packet-manager.cc:161-168) */
          a->last_data_dst = p->dst;
          if (p->next_hop == x && p->valable) agent_data_step (s, m, p, l);
          break;
        case PO_BIG_SIGN:   /* big-signaling-packet-manager.cc:97-112 */
          if (p->src != x && p->dst == x && p->next_hop == x) agent_control_step (s, m, p);
          break;
        case PO_SMALL_SIGN: /* small-signaling-packet-manager.cc:92-101 */
          if (p->dst == x && p->next_hop == x) agent_control_step (s, m, p);
          break;
        case PO_PING_FWD:   /* answered by every overlay node it crosses: packet-routing-gym.cc:254-256 */
          ping_forward_rx (s, a, p, l);
          break;
        case PO_PING_BACK:
          ping_back_rx (s, a, p);
          break;
        }
    }
  /* --- statistics are taken after the callbacks return (point-to-point-net-device.cc:432-463);
   *     the IP forward below happens inside m_rxCallback, i.e. before them, but touches none of
   *     these counters unless it drops, so the order is kept: forward first. --- */
  int consumed = 0;
  po_pkt *fwd = 0;
  if (p->first_hop)
    consumed = 1;              /* addressed to its own switch: local delivery (quirk Q6) */
  else if (p->ipdst == x)
    {
      consumed = 1;            /* leaves through device 1 to the traffic node's sink */
      /* the traffic node's own Receive counts it once more as "injected" (quirk Q3) */
      if (p->type == PO_DATA && p->last_hop == PO_LASTHOP_NONE && !p->valable
          && !(p->rejected == 1 && s->N + x < 23))
        s->c.underlay_injected++;
    }
  else
    fwd = pkt_clone (p);
  if (fwd)
    dev_send (s, s->route[x * s->N + p->ipdst], fwd);
  (void) consumed;

  if (p->type == PO_DATA && p->dst == x)
    {
      float dly = (float) (po_ns_to_seconds (s->now) - (double) p->start);
      if (p->valable)
        {
          if (p->next_hop == p->dst)
            {
              s->c.overlay_arrived++;
              push_f (&s->c.sum_cost, &s->c.n_cost, dly);
              push_f (&s->c.sum_e2e, &s->c.n_e2e, dly);
            }
        }
      else
        {
          s->c.underlay_arrived++;
          push_f (&s->c.sum_underlay_cost, &s->c.n_underlay_cost, dly);
          push_f (&s->c.sum_underlay_e2e, &s->c.n_underlay_e2e, dly);
        }
    }
  if (p->type > 0 && p->dst == x)
    s->c.bytes_signaling += p->size - 2;
  if (p->type == PO_DATA && p->last_hop == PO_LASTHOP_NONE)
    {
      if (p->valable)
        {
          s->c.overlay_injected++;
          s->c.bytes_data += p->size - 2;
        }
      else
        s->c.underlay_injected++;
    }
  free (p);
}

static void schedule_flow (po_sim *s, int f, int first);

/* PoissonGeneratorApplication::SendPacket — poisson-application.cc:297-358 */
static void
inject (po_sim *s, int f)
{
  po_flow *fl = &s->flows[f];
  po_pkt *p = (po_pkt *) calloc (1, sizeof (po_pkt));
  if (fl->big)
    {
      /* BigSignalingGeneratorApplication::SendPacket — big-signaling-application.cc:261-306: one 512-byte
       * segment of the NN weights, addressed to the neighbour's traffic node, never a routed-by-agent packet */
      p->type = PO_BIG_SIGN;
      p->valable = 0; p->rejected = 0;
      p->dst = fl->dst; p->src = fl->src; p->next_hop = fl->dst; p->last_hop = PO_LASTHOP_NONE;
      p->seg_index = fl->seg_index; p->nn_index = fl->nn_index;
      p->uid = -1;
      p->birth_ns = s->now;
      p->size = 512 + 8 + 20 + 2;
      p->first_hop = 0;
      p->ipdst = fl->dst;
      dev_send (s, s->L + fl->src, p);
      schedule_flow (s, f, 0);
      return;
    }
  p->type = PO_DATA;
  p->dst = fl->dst; p->src = fl->src; p->next_hop = fl->src; p->last_hop = PO_LASTHOP_NONE;
  p->start = (uint64_t) po_ns_to_seconds (s->now);
  p->birth_ns = s->now;
  p->size = (s->cfg.packet_size < 200 ? 200 : s->cfg.packet_size) + 8 + 20 + 2;
  int both = s->map_overlay[fl->src] >= 0 && s->map_overlay[fl->dst] >= 0;
  if (s->cfg.inject_mode == PO_INJECT_TRACE)
    {
      int64_t i = fl->trace_pos;
      p->uid = s->tr_uid[i];
      p->valable = s->tr_val[i];
      p->rejected = s->tr_rej[i];
    }
  else
    {
      uint32_t ctr[4] = { (uint32_t) s->cfg.env_id, (uint32_t) (s->cfg.env_id >> 32),
                          (uint32_t) (fl->src * s->N + fl->dst), (uint32_t) (fl->n_draw - 1) };
      uint32_t key[2] = { (uint32_t) s->cfg.seed, (uint32_t) (s->cfg.seed >> 32) };
      uint32_t r[4];
      po_philox4x32 (ctr, key, r);
      double u = ((double) r[2] + 0.5) * (1.0 / 4294967296.0);
      p->uid = s->next_uid++;
      p->valable = both ? 1 : 0;    /* U(0,1) < OverlayMask in {0,1} — sim.cc:615-617, poisson-application.cc:315-321 */
      double prej = (s->cfg.opt_agent && s->opt_rejected) ? s->opt_rejected[fl->src * s->N + fl->dst] : 0.0;
      p->rejected = u < prej ? 1 : 0;
    }
  p->first_hop = p->valable ? 1 : 0;
  p->ipdst = fl->dst;
  dev_send (s, s->L + fl->src, p);
  schedule_flow (s, f, 0);
}

/* PoissonGeneratorApplication::ScheduleNextTx — poisson-application.cc:265-295 (and the start
 * offset of sim.cc:610-631 when first != 0) */
static void
schedule_flow (po_sim *s, int f, int first)
{
  po_flow *fl = &s->flows[f];
  if (fl->big)
    {
      /* BigSignalingGeneratorApplication::ScheduleNextTx — big-signaling-application.cc:236-258
       * (StartApplication :223-232 zeroes both indices and calls it at AppStartTime, sim.cc:644) */
      fl->seg_index++;
      if (fl->seg_index >= s->big_segs) { fl->nn_index++; fl->seg_index = 0; }
    }
  if (s->cfg.inject_mode == PO_INJECT_TRACE)
    {
      int64_t i = first ? fl->trace_pos : s->tr_next[fl->trace_pos];
      fl->trace_pos = i;
      if (i >= 0) heap_push (s, s->tr_t[i], EVK_INJ, f, 0);
      return;
    }
  if (fl->big)
    {
      heap_push (s, (first ? s->big_start_ns : s->now) + s->big_gap_ns, EVK_INJ, f, 0);
      return;
    }
  uint32_t key[2] = { (uint32_t) s->cfg.seed, (uint32_t) (s->cfg.seed >> 32) };
  uint32_t r[4];
  int64_t base = s->now;
  if (first)
    {
      uint32_t ctr[4] = { (uint32_t) s->cfg.env_id, (uint32_t) (s->cfg.env_id >> 32),
                          (uint32_t) (fl->src * s->N + fl->dst), 0u };
      po_philox4x32 (ctr, key, r);
      double u0 = po_u01_53 (r[0], r[1]);
      base = 100000 + (int64_t) (u0 * 1e9);
      fl->n_draw = 1;
    }
  uint32_t ctr[4] = { (uint32_t) s->cfg.env_id, (uint32_t) (s->cfg.env_id >> 32),
                      (uint32_t) (fl->src * s->N + fl->dst), (uint32_t) fl->n_draw };
  po_philox4x32 (ctr, key, r);
  fl->n_draw++;
  double u = po_u01_53 (r[0], r[1]);
  double delay = fl->mean_s * po_neglog (u);
  int64_t t = base + (int64_t) (delay * 1e9);
  heap_push (s, t, EVK_INJ, f, 0);
}

static void
start_sim (po_sim *s)
{
  s->started = 1;
  /* ping timers are scheduled while the envs are built (sim.cc:545, data-packet-manager.cc:117-122),
   * Simulator::Stop right before Run (sim.cc:703): they get the lowest event uids. */
  for (int m = 0; m < s->M; m++)
    heap_push (s, s->cfg.ping_interval_ns, EVK_PING, m, 0);
  s->stop_seq = s->seq++;
  for (int f = 0; f < s->F; f++)
    schedule_flow (s, f, 1);
}

int32_t
po_run (po_sim *s, int64_t until_ns)
{
  if (!s->started) start_sim (s);
  int64_t stop = s->cfg.sim_time_ns;
  while (s->heap_n > 0)
    {
      po_hev *top = &s->heap[0];
      if (top->ts > stop || (top->ts == stop && top->seq > s->stop_seq))
        {
          s->now = stop;
          break;
        }
      if (top->ts >= until_ns && until_ns < stop) break;
      po_hev e = heap_pop (s);
      s->now = e.ts;
      s->c.events++;
      switch (e.kind)
        {
        case EVK_INJ: inject (s, e.a); break;
        case EVK_TXC: transmit_complete (s, e.a); break;
        case EVK_RCV: dev_receive (s, e.a, e.pkt); break;
        case EVK_PING: ping_timer (s, e.a); break;
        }
    }
  if (until_ns < stop && s->now < until_ns) s->now = until_ns;
  s->c.end_ns = s->now;
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * Construction (sim.cc:274-652) and accessors
 * ---------------------------------------------------------------------------------------- */

po_sim *
po_create (const po_config *cfg)
{
  po_sim *s = (po_sim *) calloc (1, sizeof (po_sim));
  s->cfg = *cfg;
  int N = s->N = cfg->n_nodes, M = s->M = cfg->n_overlay;
  s->L = po_num_links (N, cfg->phys_adj);
  s->links = (po_link *) calloc (s->L + N, sizeof (po_link));
  int32_t *from = (int32_t *) malloc (sizeof (int32_t) * (s->L + 1));
  int32_t *to = (int32_t *) malloc (sizeof (int32_t) * (s->L + 1));
  po_build_links (N, cfg->phys_adj, from, to);
  s->degree = (int32_t *) calloc (N, sizeof (int32_t));
  for (int i = 0; i < N; i++)
    for (int j = 0; j < N; j++)
      if (cfg->phys_adj[i * N + j] == 1) s->degree[i]++;          /* sim.cc:346-356 */
  for (int l = 0; l < s->L; l++)
    {
      po_link *k = &s->links[l];
      k->from = from[l]; k->to = to[l];
      k->bps = cfg->link_rate_bps; k->delay_ns = cfg->link_delay_ns;
      k->byte_mode = 1; k->max_size = cfg->max_out_buffer_size;   /* sim.cc:422-424 */
    }
  for (int i = 0; i < N; i++)
    {
      po_link *k = &s->links[s->L + i];
      k->from = N + i; k->to = i;
      k->bps = 1000000ull * cfg->link_rate_bps * (uint64_t) s->degree[i];   /* sim.cc:403 */
      k->delay_ns = 0; k->byte_mode = 0; k->max_size = 1000;              /* sim.cc:404-405 */
    }
  free (from); free (to);
  s->route = (int32_t *) malloc (sizeof (int32_t) * N * N);
  if (cfg->route_out) memcpy (s->route, cfg->route_out, sizeof (int32_t) * N * N);
  else po_global_routes (N, cfg->phys_adj, s->route);
  s->map_overlay = (int32_t *) malloc (sizeof (int32_t) * N);
  memcpy (s->map_overlay, cfg->map_overlay, sizeof (int32_t) * N);
  s->o2u = (int32_t *) calloc (M, sizeof (int32_t));
  for (int i = 0; i < N; i++)
    if (s->map_overlay[i] >= 0) s->o2u[s->map_overlay[i]] = i;     /* sim.cc:462-467 */
  s->agents = (po_agent *) calloc (M, sizeof (po_agent));
  for (int m = 0; m < M; m++) { s->agents[m].node = s->o2u[m]; s->agents[m].last_data_dst = -1; }
  for (int i = 0; i < M; i++)                                      /* sim.cc:469-476 */
    for (int j = i; j < M; j++)
      if (cfg->overlay_adj[i * M + j] == 1)
        {
          po_agent *a = &s->agents[i], *b = &s->agents[j];
          if (a->k < PO_MAX_K) a->nbr[a->k++] = s->o2u[j];
          if (b->k < PO_MAX_K) b->nbr[b->k++] = s->o2u[i];
        }
  for (int m = 0; m < M; m++)
    {
      int deg = 0;
      for (int j = 0; j < M; j++) if (cfg->overlay_adj[m * M + j] == 1) deg++;
      uint32_t sz = 0;                                             /* sim.cc:372-393 */
      if (cfg->signaling_type == 1) sz = 8 + 8 * (deg + 1);
      else if (cfg->signaling_type == 2) sz = 24;
      else if (cfg->signaling_type == 3) sz = 8 + 8 * (2 * deg + 1);
      s->agents[m].sign_size = sz;
    }
  if (cfg->opt_rejected)
    {
      s->opt_rejected = (double *) malloc (sizeof (double) * N * N);
      memcpy (s->opt_rejected, cfg->opt_rejected, sizeof (double) * N * N);
    }
  /* flows in creation order — sim.cc:599-632 */
  s->flows = (po_flow *) calloc (N * N, sizeof (po_flow));
  s->flow_of = (int32_t *) malloc (sizeof (int32_t) * N * N);
  uint32_t pkt = cfg->packet_size < 200 ? 200 : cfg->packet_size;  /* poisson-application.cc:277-278 */
  for (int i = 0; i < N; i++)
    for (int j = 0; j < N; j++)
      {
        s->flow_of[i * N + j] = -1;
        int both = s->map_overlay[i] >= 0 && s->map_overlay[j] >= 0;
        int active = both || cfg->activate_underlay_traffic;
        if (i != j && active && cfg->traffic_bps[i * N + j] > 0)
          {
            po_flow *fl = &s->flows[s->F];
            fl->src = i; fl->dst = j; fl->bps = cfg->traffic_bps[i * N + j];
            fl->mean_s = (double) (pkt * 8) / (double) fl->bps;
            fl->trace_pos = -1;
            s->flow_of[i * N + j] = s->F++;
          }
      }
  /* NN-weight segment sources, one per directed overlay edge that also carries traffic — sim.cc:634-647.
   * They are kept after the Poisson sources; among themselves they keep the (i, j) creation order, which is
   * the order of their simultaneous SendPacket events. */
  s->big_flow_of = (int32_t *) malloc (sizeof (int32_t) * N * N);
  for (int i = 0; i < N * N; i++) s->big_flow_of[i] = -1;
  if (cfg->train && cfg->signaling && cfg->signaling_type == 1 && cfg->big_signaling_size >= 512 && cfg->sync_step > 0)
    {
      /* big-signaling-application.cc:247-250: uint32 * 8 / float -> float, widened; uint32 512*8 / double */
      float sync = (float) cfg->sync_step;
      double rate = (double) ((float) (uint32_t) (cfg->big_signaling_size * 8u) / sync);
      s->big_gap_ns = po_seconds_to_ns ((double) (512u * 8u) / rate);
      s->big_start_ns = po_seconds_to_ns (0.0001);                 /* AppStartTime, sim.cc:244,644 */
      s->big_segs = cfg->big_signaling_size / 512u;                /* :243 */
      int nf = s->F;
      s->flows = (po_flow *) realloc (s->flows, sizeof (po_flow) * (size_t) (2 * N * N));
      memset (s->flows + nf, 0, sizeof (po_flow) * (size_t) (2 * N * N - nf));
      for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++)
          {
            int mi = s->map_overlay[i], mj = s->map_overlay[j];
            if (s->flow_of[i * N + j] < 0 || mi < 0 || mj < 0 || cfg->overlay_adj[mi * M + mj] != 1) continue;
            po_flow *fl = &s->flows[s->F];
            fl->src = i; fl->dst = j; fl->big = 1; fl->trace_pos = -1;
            s->big_flow_of[i * N + j] = s->F++;
          }
    }
  s->policy = PO_POLICY_ZERO;
  return s;
}

void
po_destroy (po_sim *s)
{
  if (!s) return;
  for (int64_t i = 0; i < s->heap_n; i++) free (s->heap[i].pkt);
  for (int l = 0; l < s->L + s->N; l++)
    for (po_pkt *p = s->links[l].q_head; p;) { po_pkt *n = p->next; free (p); p = n; }
  for (int m = 0; m < s->M; m++)
    {
      for (int t = 0; t < PO_MAX_K; t++) { free (s->agents[m].sent[t]); free (s->agents[m].win[t]); }
      free (s->agents[m].lost);
    }
  free (s->links); free (s->route); free (s->map_overlay); free (s->o2u); free (s->degree);
  free (s->agents); free (s->flows); free (s->flow_of); free (s->big_flow_of); free (s->opt_rejected); free (s->heap);
  free (s->tr_t); free (s->tr_src); free (s->tr_dst); free (s->tr_uid); free (s->tr_val); free (s->tr_rej); free (s->tr_next);
  free (s->pol_data); free (s->ev); free (s->dec);
  free (s);
}

void po_set_event_log (po_sim *s, int32_t enable) { s->log_events = enable; }

int32_t
po_set_inject_trace (po_sim *s, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst,
                     const int64_t *uid, const uint8_t *valable, const uint8_t *rejected)
{
  return po_set_inject_trace_typed (s, n, t_ns, src, dst, uid, valable, rejected, 0);
}

int32_t
po_set_inject_trace_typed (po_sim *s, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst,
                           const int64_t *uid, const uint8_t *valable, const uint8_t *rejected, const uint8_t *ptype)
{
  s->tr_n = n;
  s->tr_t = (int64_t *) malloc (sizeof (int64_t) * (n + 1)); memcpy (s->tr_t, t_ns, sizeof (int64_t) * n);
  s->tr_src = (int32_t *) malloc (sizeof (int32_t) * (n + 1)); memcpy (s->tr_src, src, sizeof (int32_t) * n);
  s->tr_dst = (int32_t *) malloc (sizeof (int32_t) * (n + 1)); memcpy (s->tr_dst, dst, sizeof (int32_t) * n);
  s->tr_uid = (int64_t *) malloc (sizeof (int64_t) * (n + 1)); memcpy (s->tr_uid, uid, sizeof (int64_t) * n);
  s->tr_val = (uint8_t *) malloc (n + 1); memcpy (s->tr_val, valable, n);
  s->tr_rej = (uint8_t *) malloc (n + 1); memcpy (s->tr_rej, rejected, n);
  s->tr_next = (int64_t *) malloc (sizeof (int64_t) * (n + 1));
  /* per-flow chains: each SendPacket schedules the flow's next one (poisson-application.cc:357) */
  int NN = s->N * s->N;
  int64_t *last = (int64_t *) malloc (sizeof (int64_t) * 2 * NN);
  for (int i = 0; i < 2 * NN; i++) last[i] = -1;
  for (int64_t i = 0; i < n; i++)
    {
      s->tr_next[i] = -1;
      int big = ptype && ptype[i] == PO_BIG_SIGN;
      int key = src[i] * s->N + dst[i];
      int f = big ? s->big_flow_of[key] : s->flow_of[key];
      if (f < 0 || (ptype && ptype[i] != PO_DATA && !big)) { free (last); return -1; }
      key += big ? NN : 0;
      if (last[key] >= 0) s->tr_next[last[key]] = i; else s->flows[f].trace_pos = i;
      last[key] = i;
    }
  free (last);
  s->cfg.inject_mode = PO_INJECT_TRACE;
  return 0;
}

int32_t
po_set_policy (po_sim *s, int32_t kind, const int32_t *data, int64_t n, uint64_t seed)
{
  s->policy = kind;
  free (s->pol_data); s->pol_data = 0; s->pol_n = n; s->pol_pos = 0;
  if (n > 0 && data)
    {
      s->pol_data = (int32_t *) malloc (sizeof (int32_t) * n);
      memcpy (s->pol_data, data, sizeof (int32_t) * n);
    }
  s->pol_rng = seed * 0x9E3779B97F4A7C15ULL + 0x1234567ULL;
  if (s->pol_rng == 0) s->pol_rng = 1;
  return 0;
}

void
po_get_counters (po_sim *s, po_counters *out)
{
  po_counters *c = &s->c;
  /* ComputeStats::getAverage: float sum / size (compute-stats-v2.cc:87-97) */
  c->avg_e2e = c->n_e2e ? c->sum_e2e / c->n_e2e : 0.0f;
  c->avg_cost = c->n_cost ? c->sum_cost / c->n_cost : 0.0f;
  c->avg_underlay_e2e = c->n_underlay_e2e ? c->sum_underlay_e2e / c->n_underlay_e2e : 0.0f;
  c->avg_underlay_cost = c->n_underlay_cost ? c->sum_underlay_cost / c->n_underlay_cost : 0.0f;
  /* ComputeStats::getSignalingOverhead — compute-stats-v2.cc:263-266 */
  c->signaling_overhead = c->bytes_data ? (float) (int32_t) c->bytes_signaling / (float) (int32_t) c->bytes_data : 0.0f;
  *out = *c;
}

int64_t po_num_events (po_sim *s) { return s->ev_n; }
const po_event *po_events (po_sim *s) { return s->ev; }
int64_t po_num_decisions (po_sim *s) { return s->dec_n; }
const po_decision *po_decisions (po_sim *s) { return s->dec; }
int32_t po_queue_bytes (po_sim *s, int32_t link) { return (int32_t) s->links[link].q_bytes; }
int32_t po_num_links_of (po_sim *s) { return s->L; }
const int32_t *po_route_table (po_sim *s) { return s->route; }
