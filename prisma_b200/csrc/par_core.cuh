/* prisma_b200 — node-parallel engine: one lane per network node, conservative lookahead inside an env.
 *
 * The serial engine (sim_core.cuh) executes one event of an environment at a time in ns-3's global
 * (timestamp, scheduler uid) order.  This engine advances the NODES of one environment concurrently
 * (a team of T lanes, lane x owns node x, its traffic node N+x, every device of both and the flows that
 * start at x) and is used by prb_run for the agents that are a function of the packet alone (next-hop
 * table).  Same physics, same reference citations as sim_core.cuh; what changes is the order in which
 * independent events are executed.
 *
 * Why that is allowed.  A node is only ever influenced from outside through packets that arrive on its
 * in-links, and a packet whose transmission starts at time s on link y->x arrives at s + tx + delay >=
 * s + LA, LA = link_delay + tx(smallest packet) (point-to-point-net-device.cc:272-302,
 * point-to-point-channel.cc:86-107).  With t_next[y] the timestamp of the earliest pending event of node
 * y and gmin their minimum, node x may execute its earliest event e iff
 *     ts(e) <  max(busy_until[y->x], min(t_next[y], gmin + LA)) + LA   for every in-link
 * (the next event y executes is one it holds already, or an arrival nobody has sent yet: not before
 * gmin + LA; strictly: an arrival with the same timestamp could still be ordered before e).  The node that holds
 * gmin always qualifies, so every round makes progress.
 *
 * Order of simultaneous events of ONE node.  ns-3 orders equal timestamps by scheduler uid, i.e. by the
 * order in which they were scheduled (scheduler.h:153-169).  That order is reproduced from local
 * information: first the time at which each event was scheduled (the transmission start of an arrival or
 * of a TransmitComplete, the previous injection of a flow, the previous ping round); when those agree too,
 * by the rules that cover the systematic cases (ping rounds: every ping timer fires at the same instant,
 * in overlay-index order, and sends tunnel by tunnel, data-packet-manager.cc:349-357; TransmitComplete is
 * scheduled before the peer's Receive, point-to-point-net-device.cc:295-296).  A tie that none of these
 * decides sets PRB_ERR_TIE: the kernel abandons the environment and the host has the event-serial engine
 * replay it (prisma_b200.cu: resolve_ties).  It takes lock-step chains longer than the recorded history
 * (full queues after a ping round) or two unrelated events that coincide to the nanosecond twice in a row.
 *
 * float32 statistics.  The reference re-sums float vectors front to back (compute-stats-v2.cc:87-97), so
 * the sums depend on the global order of deliveries and drops.  Nodes append (ts, value) records to a small
 * ring; every round the records older than gmin are folded into the sums in timestamp order by one lane
 * at a time.  Deliveries at DIFFERENT nodes in the same nanosecond are folded in node order.
 *
 * Like sim_core.cuh this header also compiles with g++ for tests/host_harness, where the lanes of a team
 * are emulated one after the other (in a shuffled order, to show that a round does not depend on it).
 */
#pragma once

#include "sim_core.cuh"

/* Row length of the link tables and wire slots per link: compile-time constants in the specialised engine
 * (PSim<LOG, DEG, W> with DEG, W > 0: index arithmetic and masks become immediates, 12.25 -> 12.05 ms per simulated
 * second on Abilene), the fields of PStatic otherwise. */
#define P_DEG (CDEG > 0 ? CDEG : ps.DEG)
#define P_W (CW > 0 ? CW : ps.W)
#ifndef PRB_SELECT_UNROLL
#define PRB_SELECT_UNROLL
#endif
#ifndef PRB_RESCAN_UNROLL
#define PRB_RESCAN_UNROLL
#endif
#ifndef PRB_SAFE_LEVEL
#define PRB_SAFE_LEVEL 1
#endif
#ifndef PRB_UNLIKELY
#define PRB_UNLIKELY(x) __builtin_expect (!!(x), 0)
#endif

namespace prb {

#ifndef PRB_P_SR
#define PRB_P_SR 8
#endif
constexpr int P_SR = PRB_P_SR;                  /* statistics-ring records per node */
constexpr int64_t P_FAR = INT64_MAX;
constexpr uint32_t P_FAR32 = 0xFFFFFFFFu;   /* no event: times inside a run are 32-bit offsets from PScal::base */
constexpr uint32_t P_NONE16 = 0xFFFFu;

/* Lineage of an event, for the order of simultaneous events (tie_before).
 * An event's scheduler uid is the order of its Schedule call, i.e. (timestamp of the parent event, uid of the parent,
 * position inside the parent's handler), recursively (scheduler.h:153-169).  Every Receive and every TransmitComplete
 * is scheduled by a transmission start (point-to-point-net-device.cc:272-302), and a transmission start happens inside
 * one event X: the Receive, injection or ping timer that called Send on an idle device, or the TransmitComplete that
 * took the next packet from the queue.  The ring record of a packet therefore doubles as the node of a lineage tree:
 * it keeps the time s of its transmission start and a reference to X (kind, link / flow / timer, ring index or
 * ordinal, position of the Send inside X's handler).  Two simultaneous events are ordered by walking both lineages
 * upwards, parent by parent, until the parents' timestamps differ (the earlier parent ran first) or the parents are
 * one and the same event (the Schedule call that came first inside its handler wins); source and ping-timer events
 * recompute their parents (previous injection: Philox / trace; previous firing: one ping interval earlier; every ping
 * timer fires at the same instants in overlay-index order, sim.cc:519-560).  Records stay readable in their ring
 * after they have been consumed, until the ring wraps (PStatic: rings are sized for the live packets plus a lineage
 * retention); a walk that meets an overwritten record cannot decide and sets PRB_ERR_TIE.  Nothing of this is touched
 * by an event that has no simultaneous rival: the hot path only stores (s, X) with the record it writes anyway. */
enum { PK_NONE = 0, PK_RCV = 1, PK_TXC = 2, PK_FLOW = 3, PK_PING = 4 };
enum { PS_OVERLAY_ARRIVED = 0, PS_UNDERLAY_ARRIVED = 1, PS_OVERLAY_DROP = 2, PS_UNDERLAY_DROP = 3 };

/* packet record inside a link ring — 32 bytes (one sector) */
struct PRec
{
  uint64_t w0;       /* same packing as Pkt::w0 */
  uint32_t uid;
  uint32_t aux;
  int64_t s;         /* absolute time of the transmission start that carries the packet (written when it starts) */
  uint64_t xw;       /* the event that started it: kind 3 | position 5 | idx 16 | seq 24 | own ring index 16 */
};

/* reference to an event: RCV / TXC (idx = link, seq = ring index of the record), FLOW (idx = flow, seq = f_pos at the
 * injection), PING (idx = overlay index, seq = round); t = its timestamp */
struct PEv
{
  int32_t kind, idx;
  uint32_t seq;
  int64_t t;
};

constexpr uint64_t P_XW_QUEUED = ~0ull;      /* a record that waits in the queue: no transmission start yet */
constexpr int P_POS_LAST = 255;              /* the Schedule call a handler makes after all its Sends */
constexpr int P_WALK_MAX = 8192;             /* generations a lineage walk may climb */

PRB_HD uint64_t
xw_pack (int kind, uint32_t pos, int idx, uint32_t seq, uint32_t self)
{
  return ((uint64_t) (uint32_t) kind << 61) | ((uint64_t) (pos & 31u) << 56) | ((uint64_t) ((uint32_t) idx & 0xFFFFu) << 40)
         | ((uint64_t) (seq & 0xFFFFFFu) << 16) | (uint64_t) (self & 0xFFFFu);
}

struct PStat
{
  uint32_t ts;       /* offset from PScal::base */
  uint32_t code;
  float v;
};

struct PStatic
{
  int32_t W;                 /* wire-ring slots per link (power of two) */
  int32_t DEG;               /* row length of in_links / out_links */
  int32_t ring_records;      /* PRec per environment */
  int32_t walk_max;          /* generations a lineage walk may climb (P_WALK_MAX; PRB_LINEAGE_WALK_MAX for tests) */
  uint32_t la32;             /* la_ns as a 32-bit offset */
  int64_t la_ns;             /* smallest (tx + delay) of a switch<->switch link */
  int64_t max_chunk_ns;      /* longest stretch of simulated time one launch may cover (32-bit offsets) */
  const int16_t *in_links;   /* [N*DEG] switch links that end at x, -1 padded */
  const int16_t *out_links;  /* [N*DEG] switch links that start at x, -1 padded */
  const int32_t *ring_off;   /* [LA] first record of the ring of link l */
  const uint16_t *ring_mask; /* [LA] ring capacity - 1 */
  const int32_t *flow_begin; /* [N+1] flows with source x: [flow_begin[x], flow_begin[x+1]) */
  const int32_t *tr_prev;    /* trace mode: previous injection of the same flow, -1 for the first */
  const int32_t *pol_off;    /* per-node action trace: [M+1] offsets into pol_data */
  int32_t off_scal, off_t_next, off_t_ping, off_fmin_ts, off_f_ts, off_busy, off_w_ts, off_sring, off_q_bytes, off_f_pos,
      off_ping_index, off_pol_pos, off_fmin_f, off_head, off_mid, off_tail, off_s_head, off_s_tail, off_fly_szc, off_cand;
  int32_t pblob_bytes;
};

struct PScal
{
  int64_t front_ts;          /* every event with ts < front_ts has executed */
  int64_t base;              /* absolute time of offset 0 of the 32-bit times in this blob */
  int64_t ev_n;
  int64_t tx_started;        /* TransmitStart calls so far (each schedules one TransmitComplete) */
  int64_t ev_real;           /* Receive / injection / ping-timer events executed */
  int32_t started, finished;
  prb_counters c;
};

struct PDyn
{
  PScal *sc;
  int64_t *t_ping, *f_ts;                   /* absolute ns */
  uint32_t *t_next, *fmin_ts, *busy, *w_ts;   /* offsets from sc->base; busy: offset + 1, 0 = idle since before base */
  PStat *sring;
  uint32_t *q_bytes, *f_pos, *ping_index, *pol_pos;
  int32_t *fmin_f, *cand;
  uint16_t *head, *mid, *tail;
  uint8_t *s_head, *s_tail, *fly_szc;
  PRec *rings;
  prb_event *ev_log;
};

PRB_HD void
p_bind_state (const PStatic &ps, char *blob, PDyn &d)
{
  d.sc = (PScal *) (blob + ps.off_scal);
  d.t_next = (uint32_t *) (blob + ps.off_t_next);
  d.t_ping = (int64_t *) (blob + ps.off_t_ping);
  d.fmin_ts = (uint32_t *) (blob + ps.off_fmin_ts);
  d.f_ts = (int64_t *) (blob + ps.off_f_ts);
  d.busy = (uint32_t *) (blob + ps.off_busy);
  d.w_ts = (uint32_t *) (blob + ps.off_w_ts);
  d.sring = (PStat *) (blob + ps.off_sring);
  d.q_bytes = (uint32_t *) (blob + ps.off_q_bytes);
  d.f_pos = (uint32_t *) (blob + ps.off_f_pos);
  d.ping_index = (uint32_t *) (blob + ps.off_ping_index);
  d.pol_pos = (uint32_t *) (blob + ps.off_pol_pos);
  d.fmin_f = (int32_t *) (blob + ps.off_fmin_f);
  d.cand = (int32_t *) (blob + ps.off_cand);
  d.head = (uint16_t *) (blob + ps.off_head);
  d.mid = (uint16_t *) (blob + ps.off_mid);
  d.tail = (uint16_t *) (blob + ps.off_tail);
  d.s_head = (uint8_t *) (blob + ps.off_s_head);
  d.s_tail = (uint8_t *) (blob + ps.off_s_tail);
  d.fly_szc = (uint8_t *) (blob + ps.off_fly_szc);
}

/* per-lane counters of one launch (registers on the device) */
struct PAcc
{
  uint32_t overlay_injected, overlay_arrived, overlay_lost, underlay_injected, underlay_arrived, underlay_lost;
  uint32_t bytes_data, bytes_signaling, packet_hops, data_hops, env_steps, decisions, vanished, ev_real, tx_started;
  uint32_t errors;
};

struct PCand
{
  uint32_t ts;      /* offset from sc->base */
  int32_t kind;
  int32_t idx;      /* link (RCV, TXC), flow, overlay index */
};

PRB_HD uint32_t
sat32 (uint32_t a, uint32_t b)
{
  return a > P_FAR32 - b ? P_FAR32 : a + b;
}

/* absolute ns -> offset from base, saturated */
PRB_HD uint32_t
rel32 (int64_t t, int64_t base)
{
  if (t == P_FAR) return P_FAR32;
  uint64_t dlt = (uint64_t) (t - base);
  return dlt > 0xFFFFFFFEull ? P_FAR32 : (uint32_t) dlt;
}

/* the running event as a parent: when it was scheduled, and which event it is */
struct PCtx
{
  int64_t sched;     /* P_SCHED_LAZY: an injection, whose scheduling time is recomputed only if somebody asks */
  int32_t kind, idx; /* PEv fields of the running event */
  uint32_t seq;
};
constexpr int64_t P_SCHED_LAZY = INT64_MIN + 1;

/* gap (nd >= 1) or start offset (nd == 0) of synthetic source f of environment env: Philox-4x32-10, DESIGN.md section 2 */
PRB_HDN inline int64_t
p_philox_gap (const Static &st, int env, int f, uint32_t nd)
{
  uint64_t eid = st.env_id_base + (uint64_t) env;
  uint32_t fkey = (uint32_t) (st.flow_src[f] * st.N + st.flow_dst[f]);
  uint32_t r[4];
  philox4x32 ((uint32_t) eid, (uint32_t) (eid >> 32), fkey, nd, (uint32_t) st.seed, (uint32_t) (st.seed >> 32), r);
  if (nd == 0) return 100000 + (int64_t) dmul_ (u01_53 (r[0], r[1]), 1e9);      /* start offset of the source */
  double u = u01_53 (r[0], r[1]);
  double delay = dmul_ (st.flow_mean_s[f], neglog (u));
  return (int64_t) dmul_ (delay, 1e9);
}

/* ---- lineage (see PRec) ---- */
/* The event that scheduled e.  Returns 1 and fills p (with its timestamp) and pos (the place of the Schedule call
 * inside p's handler: Send number * 2, + 1 for the peer's Receive, which TransmitStart schedules after the
 * TransmitComplete, point-to-point-net-device.cc:295-296; P_POS_LAST for what a source or a ping timer schedules after
 * its Sends, poisson-application.cc:331,357); 2 when e was scheduled before Simulator::Run (p.idx = creation rank: ping
 * timers in overlay-index order, then the sources, sim.cc:519-560,599-632); 0 when the record has been overwritten. */
PRB_HD int
p_parent_of (const Static &st, const PStatic &ps, const PRec *rings, int env, const PEv &e, PEv &p, int &pos)
{
  if (e.kind == PK_RCV || e.kind == PK_TXC)
    {
      const PRec &r = rings[ps.ring_off[e.idx] + (e.seq & ps.ring_mask[e.idx])];
      const uint64_t xw = r.xw;
      const int kind = (int) (xw >> 61);
      if ((uint32_t) (xw & 0xFFFFu) != (e.seq & 0xFFFFu) || kind == PK_NONE || kind > PK_PING) return 0;
      p.kind = kind; p.idx = (int32_t) ((xw >> 40) & 0xFFFFu); p.seq = (uint32_t) ((xw >> 16) & 0xFFFFFFu); p.t = r.s;
      pos = (int) ((xw >> 56) & 31u) * 2 + (e.kind == PK_RCV ? 1 : 0);
      return 1;
    }
  pos = P_POS_LAST;
  if (e.kind == PK_FLOW)
    {
      const int f = e.idx;
      if (st.inject_mode == PRB_INJECT_TRACE)
        {
          int32_t pv = ps.tr_prev[(int32_t) e.seq];
          if (pv < 0)
            {
              p.kind = PK_NONE; p.idx = st.M + f; p.seq = 0; p.t = -1;
              return 2;
            }
          p.kind = PK_FLOW; p.idx = f; p.seq = (uint32_t) pv; p.t = st.tr_t[pv];
          return 1;
        }
      uint32_t nd = e.seq - 1;          /* ordinal of the gap that produced e.t */
      if (nd <= 1)
        {
          p.kind = PK_NONE; p.idx = st.M + f; p.seq = 0; p.t = -1;
          return 2;
        }
      p.kind = PK_FLOW; p.idx = f; p.seq = e.seq - 1; p.t = e.t - p_philox_gap (st, env, f, nd);
      return 1;
    }
  /* ping timer */
  if (e.seq == 0)
    {
      p.kind = PK_NONE; p.idx = e.idx; p.seq = 0; p.t = -1;
      return 2;
    }
  p.kind = PK_PING; p.idx = e.idx; p.seq = e.seq - 1; p.t = e.t - st.ping_interval_ns;
  return 1;
}

/* Scheduler order of two events with one timestamp: -1 a first, +1 b first, 0 undecided (a record on the way up has
 * been overwritten, or the walk is longer than P_WALK_MAX generations). */
PRB_NOINLINE int
p_lineage_cmp (const Static &st, const PStatic &ps, const PRec *rings, int env, uint64_t ak, int64_t at, uint64_t bk, int64_t bt)
{
  /* the two events travel as scalars (kind << 48 | idx << 32 | seq, timestamp): register arguments, no stack frame */
  PEv a, b;
  a.kind = (int32_t) (ak >> 48); a.idx = (int32_t) ((ak >> 32) & 0xFFFFu); a.seq = (uint32_t) ak; a.t = at;
  b.kind = (int32_t) (bk >> 48); b.idx = (int32_t) ((bk >> 32) & 0xFFFFu); b.seq = (uint32_t) bk; b.t = bt;
  for (int depth = 0; depth < ps.walk_max; depth++)
    {
      if (a.kind == b.kind && a.idx == b.idx && a.seq == b.seq) return 0;       /* one event */
      /* every ping timer fires at the same instants: two firings with one timestamp belong to one round and run in
       * the order the timers were created */
      if (a.kind == PK_PING && b.kind == PK_PING) return a.idx < b.idx ? -1 : 1;
      PEv pa, pb;
      int posa, posb;
      const int ra = p_parent_of (st, ps, rings, env, a, pa, posa), rb = p_parent_of (st, ps, rings, env, b, pb, posb);
      if (ra == 0 || rb == 0) return 0;
      if (pa.t != pb.t) return pa.t < pb.t ? -1 : 1;
      if (ra == 2 || rb == 2)
        {
          if (ra != rb) return ra == 2 ? -1 : 1;       /* cannot happen (setup time is -1); keep the order total */
          return pa.idx < pb.idx ? -1 : (pa.idx > pb.idx ? 1 : 0);
        }
      if (pa.kind == pb.kind && pa.idx == pb.idx && pa.seq == pb.seq)
        return posa < posb ? -1 : (posa > posb ? 1 : 0);
      a = pa; b = pb;
    }
  return 0;
}


template <bool LOG, int CDEG = 0, int CW = 0>
struct PSim
{
  const Static &st;
  const PStatic &ps;
  PDyn &d;
  PAcc &acc;
  int env;
  int64_t base;              /* absolute time of offset 0 (sc->base) */
  uint32_t until_rel, stop_rel;

  PRB_HD PSim (const Static &st_, const PStatic &ps_, PDyn &d_, PAcc &acc_, int env_) : st (st_), ps (ps_), d (d_), acc (acc_), env (env_), base (0), until_rel (0), stop_rel (0) {}

  /* the run covers events before until_ns; offsets are taken from the run's start (sc->base) */
  PRB_HD void
  set_window (int64_t until_ns)
  {
    base = d.sc->base;
    until_rel = rel32 (until_ns, base);
    stop_rel = rel32 (st.sim_time_ns, base);
    if (until_rel == P_FAR32) until_rel = P_FAR32 - 1;
    if (stop_rel == P_FAR32) stop_rel = P_FAR32 - 1;
    until_is_stop = until_ns >= st.sim_time_ns;
  }
  bool until_is_stop;

  PRB_HD static void
  to_pkt (const PRec &r, Pkt &p)
  {
    p.w0 = r.w0; p.prev_t = 0; p.uid = r.uid; p.aux = r.aux; p.rcv_ts = 0; p.rcv_seq = 0; p.next = NIL; p.pad = 0;
  }
  PRB_HD static void
  to_rec (const Pkt &p, PRec &r)
  {
    r.w0 = p.w0; r.uid = p.uid; r.aux = p.aux; r.s = 0; r.xw = P_XW_QUEUED;
  }

  PRB_HD void
  log_event (int64_t now, int kind, int node, int link, const Pkt &p, int qbytes)
  {
    if (!LOG) return;
    int64_t i;
#if defined(__CUDA_ARCH__)
    i = (int64_t) atomicAdd ((unsigned long long *) &d.sc->ev_n, 1ull);
#else
    i = d.sc->ev_n++;
#endif
    if (i >= st.ev_cap)
      {
        if (st.ev_cap > 0) acc.errors |= PRB_ERR_LOG;
        return;
      }
    prb_event *e = &d.ev_log[i];
    int data = p_type (p) == PRB_DATA;
    e->t_ns = now; e->kind = kind; e->node = node; e->link = link; e->pad_ = 0;
    e->uid = data ? (int64_t) p.uid : -1;
    e->size = st.szc_bytes[p_szc (p)]; e->type = p_type (p); e->src = p_src (p); e->dst = p_dst (p);
    e->last_hop = p_last_hop (p) == NODE_NONE ? PRB_LASTHOP_NONE : p_last_hop (p);
    e->next_hop = p_next_hop (p); e->valable = data ? p_valable (p) : -1;
    e->qbytes = qbytes;
  }

  /* ---- float statistics: appended here, folded in timestamp order by the team ---- */
  PRB_HD void
  stat_push (int x, uint32_t ts, int64_t sched, uint32_t code, float v)
  {
    uint8_t t = d.s_tail[x];
    PStat &s = d.sring[x * P_SR + (t & (P_SR - 1))];
    /* code in the top two bits; below, how long before ts the event was scheduled (saturated): records of
     * different nodes with one timestamp are folded earliest-scheduled first, like the events ran */
    int64_t dd = sched == P_SCHED_LAZY ? 0 : (sched < 0 ? 0x3FFFFFFF : base + ts - sched);
    s.ts = ts; s.code = (code << 30) | (uint32_t) (dd > 0x3FFFFFFF ? 0x3FFFFFFF : dd); s.v = v;
    d.s_tail[x] = (uint8_t) (t + 1);
  }
  PRB_HD int
  stat_room (int x) const
  {
    return P_SR - (int) (uint8_t) (d.s_tail[x] - d.s_head[x]);
  }

  PRB_HD void
  count_drop (int x, uint32_t now, int64_t sched, bool overlay_side)
  {
    if (overlay_side)
      acc.overlay_lost++;
    else
      acc.underlay_lost++;
    stat_push (x, now, sched, overlay_side ? PS_OVERLAY_DROP : PS_UNDERLAY_DROP, st.loss_penalty);
  }

  /* One request of the running event to a device of its node: mode 0 = PointToPointNetDevice::Send +
   * Queue::DoEnqueue, mode 1 = the TransmitComplete of link l starts the packet at the head of its queue
   * (already in the ring).  cur: the running event; pos: position of this Send among the Sends of its handler
   * (the tunnel index inside a ping round).  There is ONE call site (execute), so one copy in the instruction stream. */
  PRB_HD void
  dev_send (int x, int l, uint32_t now, const Pkt &p, const PCtx &cur, uint32_t pos, int mode)
  {
    const int64_t now_abs = base + now;
    int node = st.link_from[l];
    int data = p_type (p) == PRB_DATA;
    uint32_t size = st.szc_bytes[p_szc (p)];
    uint32_t qb = d.q_bytes[l];
    uint16_t m = d.mid[l], t = d.tail[l];
    uint16_t mask = ps.ring_mask[l];
    PRec *slot;
    bool busy = false;
    if (mode == 0)
      {
        if (PRB_UNLIKELY (data && p_rejected (p) && node < 23))
          {
            count_drop (x, now, cur.sched, p_valable (p));
            return;
          }
        if (l >= st.L) log_event (now_abs, PRB_EV_INJ, node, l, p, -1);
        bool fits = l < st.L ? (qb + size <= st.max_out_buffer) : ((uint32_t) (uint16_t) (t - m) + 1 <= 1000u);
        if (PRB_UNLIKELY (!fits))
          {
            if (l < st.L) log_event (now_abs, PRB_EV_DRP, node, l, p, (int) qb);
            if (data) count_drop (x, now, cur.sched, p_valable (p) && p_dst (p) != node);
            return;
          }
        if (l < st.L) log_event (now_abs, PRB_EV_ENQ, node, l, p, (int) (qb + size));
        if (PRB_UNLIKELY ((uint16_t) (t - d.head[l]) > mask))
          {
            acc.errors |= PRB_ERR_POOL;
            return;
          }
        /* a TransmitComplete with nothing waiting only marks the device idle: it is not an event here, the
         * device is READY as soon as the clock reaches busy_until */
        busy = m != t || d.busy[l] > now + 1;
        if (PRB_UNLIKELY (!busy && d.busy[l] == now + 1)) busy = completes_later (l, now_abs, cur);
        slot = &d.rings[ps.ring_off[l] + (t & mask)];
        d.tail[l] = (uint16_t) (t + 1);
        if (busy)
          {
            PRec r;
            to_rec (p, r);
            *slot = r;
            d.q_bytes[l] = qb + size;
            return;
          }
      }
    else
      {
        slot = &d.rings[ps.ring_off[l] + (m & mask)];
        qb -= size;
        d.q_bytes[l] = qb;
      }
    /* PointToPointNetDevice::TransmitStart + PointToPointChannel::TransmitStart; the record keeps who started it */
    const uint64_t xw = xw_pack (cur.kind, pos, cur.idx, cur.seq, m);
    if (mode == 0)
      {
        PRec r;
        to_rec (p, r);
        r.s = now_abs; r.xw = xw;
        *slot = r;
      }
    else
      {
        slot->s = now_abs; slot->xw = xw;
      }
    if (l < st.L) log_event (now_abs, PRB_EV_TXB, node, l, p, (int) qb);
    int szc = p_szc (p);
    uint32_t tx = (uint32_t) st.tx_ns[st.link_rc[l] * MAX_SZC + szc];
    uint32_t delay = l < st.L ? (uint32_t) st.link_delay_ns : 0u;
    d.busy[l] = now + tx + 1;
    d.fly_szc[l] = (uint8_t) szc;
    if ((uint16_t) (m - d.head[l]) >= (uint16_t) P_W) acc.errors |= PRB_ERR_POOL;
    d.w_ts[l * P_W + (m & (P_W - 1))] = now + tx + delay;
    d.mid[l] = (uint16_t) (m + 1);
    acc.tx_started++;
  }

  /* the pending TransmitComplete of link l: scheduled by the transmission start of the record in front of mid */
  PRB_HD PCtx
  txc_ctx (int l, int64_t ts) const
  {
    PCtx c;
    int k = st.link_rc[l] * MAX_SZC + d.fly_szc[l];
    c.sched = ts - st.tx_ns[k];
    c.kind = PK_TXC; c.idx = l; c.seq = (uint16_t) (d.mid[l] - 1);
    return c;
  }

  /* The device completes in this very nanosecond.  If that TransmitComplete comes after the running event in
   * scheduler order the device is still BUSY: the packet is queued and the completion, now a real event of
   * this node at the same timestamp, starts it. */
  PRB_HDN bool
  completes_later (int l, int64_t now, const PCtx &cur)
  {
    PCtx tc = txc_ctx (l, now);
    int r = lineage_cmp (ev_of (tc, now), ev_of (cur, now));
    if (r == 0) acc.errors |= PRB_ERR_TIE;
    return r >= 0;
  }

  PRB_HD int32_t
  policy_action (int m, const Pkt &p)
  {
    if (st.policy == PRB_POLICY_TABLE) return st.pol_data[m * st.M + st.map_overlay[p_dst (p)]];
    if (st.policy == PRB_POLICY_NODE_TRACE)
      {
        uint32_t i = d.pol_pos[m];
        if ((int64_t) ps.pol_off[m] + i >= (int64_t) ps.pol_off[m + 1])
          {
            acc.errors |= PRB_ERR_TRACE;
            return 0;
          }
        d.pol_pos[m] = i + 1;
        return st.pol_data[ps.pol_off[m] + i];
      }
    return 0;
  }

  PRB_HD static void
  blank (Pkt &q)
  {
    q.w0 = 0; q.prev_t = 0; q.rcv_ts = 0; q.rcv_seq = 0; q.pad = 0; q.next = NIL; q.uid = 0; q.aux = 0;
  }

  /* a control packet of node x back to node lh (ping echo, small signalling) */
  PRB_HD static void
  reply (Pkt &q, int type, int x, int lh, int szc)
  {
    blank (q);
    set_p_type (q, type);
    set_p_dst (q, lh); set_p_next_hop (q, lh); set_p_last_hop (q, x); set_p_src (q, x); set_p_ipdst (q, lh);
    set_p_prev_node (q, NODE_NONE);
    set_p_szc (q, szc);
  }

  /* the Receive of the record at ring index i of link l */
  PRB_HD PCtx
  rcv_ctx (int l, int64_t ts, const PRec &r, uint32_t i) const
  {
    PCtx c;
    int k = st.link_rc[l] * MAX_SZC + (int) ((r.w0 >> 7) & 15);
    c.sched = ts - st.tx_ns[k] - (l < st.L ? st.link_delay_ns : 0);
    c.kind = PK_RCV; c.idx = l; c.seq = i & 0xFFFFu;
    return c;
  }

  /* the pending firing of ping timer m (round ping_index[m]) */
  PRB_HD PCtx
  ping_ctx (int m, int64_t ts) const
  {
    PCtx c;
    c.sched = d.ping_index[m] == 0 ? -1 : ts - st.ping_interval_ns;
    c.kind = PK_PING; c.idx = m; c.seq = d.ping_index[m];
    return c;
  }

  /* ---- sources ---- */
  PRB_HD int64_t
  philox_gap (int f, uint32_t nd) const
  {
    return p_philox_gap (st, env, f, nd);
  }

  PRB_HD void
  schedule_flow (int f, bool first, int64_t now)
  {
    if (st.inject_mode == PRB_INJECT_TRACE)
      {
        int32_t i = first ? st.flow_first[f] : st.tr_next[(int32_t) d.f_pos[f]];
        d.f_pos[f] = (uint32_t) i;
        d.f_ts[f] = i >= 0 ? st.tr_t[i] : P_FAR;
        return;
      }
    int64_t base = now;
    uint32_t nd = d.f_pos[f];
    if (first)
      {
        base = philox_gap (f, 0u);
        nd = 1;
      }
    d.f_pos[f] = nd + 1;
    d.f_ts[f] = base + philox_gap (f, nd);
  }

  /* when was the pending injection of flow f scheduled: the flow's previous injection (or setup) */
  PRB_HD int64_t
  flow_sched_time (int f) const
  {
    if (st.inject_mode == PRB_INJECT_TRACE)
      {
        int32_t i = (int32_t) d.f_pos[f];
        int32_t pv = i >= 0 ? ps.tr_prev[i] : -1;
        return pv >= 0 ? st.tr_t[pv] : -1;
      }
    uint32_t nd = d.f_pos[f] - 1;       /* ordinal of the gap that produced f_ts */
    if (nd <= 1) return -1;
    return d.f_ts[f] - philox_gap (f, nd);
  }

  /* the running / pending injection of flow f as a parent; prev: time of the flow's previous injection */
  PRB_HD PCtx
  flow_ctx (int f, int64_t prev) const
  {
    PCtx c;
    c.sched = prev;
    c.kind = PK_FLOW; c.idx = f; c.seq = d.f_pos[f];
    return c;
  }

  /* ---- lineage (see PRec) ---- */
  PRB_HD static PEv
  ev_of (const PCtx &c, int64_t t)
  {
    PEv e;
    e.kind = c.kind; e.idx = c.idx; e.seq = c.seq; e.t = t;
    return e;
  }

  /* scheduler order of two events with one timestamp (p_lineage_cmp, out of line: rare) */
  PRB_HD int
  lineage_cmp (const PEv &a, const PEv &b) const
  {
    return p_lineage_cmp (st, ps, d.rings, env, ((uint64_t) (uint32_t) a.kind << 48) | ((uint64_t) ((uint32_t) a.idx & 0xFFFFu) << 32) | a.seq, a.t,
                          ((uint64_t) (uint32_t) b.kind << 48) | ((uint64_t) ((uint32_t) b.idx & 0xFFFFu) << 32) | b.seq, b.t);
  }

  /* earliest injection among the flows of node x; equal times: the one scheduled first */
  PRB_HD void
  rescan_flows (int x)
  {
    int64_t bt = P_FAR;
    int bf = -1;
    PRB_RESCAN_UNROLL
    for (int f = ps.flow_begin[x]; f < ps.flow_begin[x + 1]; f++)
      {
        int64_t t = d.f_ts[f];
        if (t < bt)
          {
            bt = t; bf = f;
          }
        else if (PRB_UNLIKELY (t == bt && t != P_FAR))
          bf = flow_tie (bf, f);
      }
    d.fmin_ts[x] = rel32 (bt, base);
    d.fmin_f[x] = bf;
  }

  PRB_HDN int
  flow_tie (int bf, int f)
  {
    int64_t a = flow_sched_time (bf), b = flow_sched_time (f);
    if (a == b && a >= 0) acc.errors |= PRB_ERR_TIE;
    return b < a ? f : bf;       /* same setup time (-1): creation order, i.e. the lower index stays */
  }

  /* ---- order of two simultaneous events of one node ---- */
  PRB_HD PCtx
  cand_ctx (const PCand &c) const
  {
    switch (c.kind)
      {
      case PK_RCV:
        return rcv_ctx (c.idx, base + c.ts, d.rings[ps.ring_off[c.idx] + (d.head[c.idx] & ps.ring_mask[c.idx])], d.head[c.idx]);
      case PK_TXC:
        return txc_ctx (c.idx, base + c.ts);
      case PK_FLOW:
        return flow_ctx (c.idx, flow_sched_time (c.idx));
      default:
        return ping_ctx (c.idx, base + c.ts);
      }
  }

  /* true when a runs before b (same timestamp, same node): the event that was scheduled earlier has the smaller
   * scheduler uid; scheduled at the same instant, the lineages decide (lineage_cmp).  A pair the walk cannot decide sets
   * PRB_ERR_TIE unless the two events cannot influence each other (TransmitComplete of two different devices). */
  PRB_HDN bool
  tie_before (const PCand &a, const PCand &b)
  {
    PCtx ca = cand_ctx (a), cb = cand_ctx (b);
    if (ca.sched != cb.sched) return ca.sched < cb.sched;
    int r = lineage_cmp (ev_of (ca, base + a.ts), ev_of (cb, base + b.ts));
    if (r != 0) return r < 0;
    if (!(a.kind == PK_TXC && b.kind == PK_TXC))
      {
#if defined(PRB_DEBUG_TIES) && !defined(__CUDA_ARCH__)
        fprintf (stderr, "tie: ts=%lld sched=%lld a=(%d,%d,%u) b=(%d,%d,%u)\n", (long long) (base + a.ts), (long long) ca.sched,
                 a.kind, a.idx, ca.seq, b.kind, b.idx, cb.seq);
#endif
        acc.errors |= PRB_ERR_TIE;
      }
    return a.kind != b.kind ? a.kind < b.kind : a.idx < b.idx;
  }

  PRB_HD void
  offer (PCand &best, uint32_t ts, int kind, int idx)
  {
    if (ts < best.ts)
      {
        best.ts = ts; best.kind = kind; best.idx = idx;
      }
    else if (PRB_UNLIKELY (ts == best.ts && ts != P_FAR32))
      {
        PCand c;
        c.ts = ts; c.kind = kind; c.idx = idx;
        if (tie_before (c, best)) best = c;
      }
  }

  /* earliest pending event of node x */
  PRB_HD PCand
  select (int x)
  {
    PCand b;
    b.ts = P_FAR32; b.kind = PK_NONE; b.idx = -1;
    const int W = P_W;
    const int la = st.L + x;
    /* the links that end at x (the access link last), then the ones that start at x */
    PRB_SELECT_UNROLL
    for (int j = 0; j <= P_DEG; j++)
      {
        int l = j < P_DEG ? ps.in_links[x * P_DEG + j] : la;
        if (l < 0) continue;
        uint16_t h = d.head[l];
        if (h != d.mid[l]) offer (b, d.w_ts[l * W + (h & (W - 1))], PK_RCV, l);
      }
    PRB_SELECT_UNROLL
    for (int j = 0; j <= P_DEG; j++)
      {
        int l = j < P_DEG ? ps.out_links[x * P_DEG + j] : la;
        if (l < 0) continue;
        if (d.mid[l] != d.tail[l]) offer (b, d.busy[l] - 1, PK_TXC, l);
      }
    int m = st.map_overlay[x];
    if (m >= 0) offer (b, rel32 (d.t_ping[m], base), PK_PING, m);
    if (d.fmin_f[x] >= 0) offer (b, d.fmin_ts[x], PK_FLOW, d.fmin_f[x]);
    return b;
  }

  /* the record the handler of c will read first (a packet at the head of a wire or of a queue): ask for its line
   * while the team still agrees on the round (select -> execute is a few thousand cycles; the record was written by
   * a lane of this warp and sits in L2) */
  PRB_HD void
  prefetch_record (const PCand &c) const
  {
#if defined(__CUDA_ARCH__) && !defined(PRB_NO_PREFETCH)
    if (c.kind == PK_RCV || c.kind == PK_TXC)
      {
        const int l = c.idx;
        uint16_t i = c.kind == PK_RCV ? d.head[l] : d.mid[l];
        const PRec *p = &d.rings[ps.ring_off[l] + (i & ps.ring_mask[l])];
        asm volatile ("prefetch.global.L1 [%0];" ::"l"(p));
      }
#endif
  }

  /* does c lie before the end of this run and of the simulation? */
  PRB_HD bool
  in_time (const PCand &c) const
  {
    if (c.kind == PK_NONE) return false;
    if (c.ts < until_rel && c.ts < stop_rel) return true;
    /* Simulator::Stop is scheduled after the first ping timers: a first ping at exactly simTime still runs */
    return c.ts == stop_rel && until_is_stop && c.kind == PK_PING && d.ping_index[c.idx] == 0;
  }

  /* Lower bounds on the time of the next event a node can still execute (t_next[] of every node is published,
   * gmin is their minimum).  Level 0: its own pending events, or an arrival nobody has sent yet (>= gmin + LA).
   * Level 1: the unsent arrivals are bounded through the node's in-links with the level-0 bounds of its
   * neighbours: a link delivers no sooner than LA after max (end of its current transmission, the sender's
   * next event). */
  PRB_HD uint32_t
  horizon0 (int y, uint32_t gmin) const
  {
    uint32_t tn = d.t_next[y], g = sat32 (gmin, ps.la32);
    return tn < g ? tn : g;
  }

  PRB_HD uint32_t
  link_bound (int l, uint32_t e_from) const
  {
    uint32_t b = d.busy[l];
    b = b ? b - 1 : 0;
#ifdef PRB_EXP_QHEAD
    if (d.mid[l] != d.tail[l])
      {
        const PRec &r = d.rings[ps.ring_off[l] + (d.mid[l] & ps.ring_mask[l])];
        uint32_t tx = (uint32_t) st.tx_ns[st.link_rc[l] * MAX_SZC + (int) ((r.w0 >> 7) & 15)];
        return sat32 (b, tx + (uint32_t) st.link_delay_ns);
      }
#endif
    return sat32 (b > e_from ? b : e_from, ps.la32);
  }

  PRB_HD uint32_t
  horizon1 (int y, uint32_t gmin) const
  {
    uint32_t e = d.t_next[y];
    for (int j = 0; j < P_DEG; j++)
      {
        int l = ps.in_links[y * P_DEG + j];
        if (l < 0) break;
        uint32_t a = link_bound (l, horizon0 (st.link_from[l], gmin));
        e = a < e ? a : e;
      }
    return e;
  }

  /* may node x execute c now? */
  PRB_HD bool
  safe (int x, const PCand &c, uint32_t gmin) const
  {
    if (!in_time (c)) return false;
    if (stat_room (x) < 2) return false;
#if PRB_SAFE_LEVEL == 0
    if (c.ts >= sat32 (gmin, 2 * ps.la32)) return false;
#endif
    for (int j = 0; j < P_DEG; j++)
      {
        int l = ps.in_links[x * P_DEG + j];
        if (l < 0) break;
#if PRB_SAFE_LEVEL == 0
        if (c.ts >= link_bound (l, d.t_next[st.link_from[l]])) return false;
#elif PRB_SAFE_LEVEL == 2
        if (c.ts >= link_bound (l, horizon1 (st.link_from[l], gmin))) return false;
#else
        if (c.ts >= link_bound (l, horizon0 (st.link_from[l], gmin))) return false;
#endif
      }
    return true;
  }

  /* Executes event c of node x.  Staged so that lanes running different kinds of events share the expensive
   * steps: (A) the ring record of a Receive or of a TransmitComplete is fetched by one load, (B) the kind's
   * own logic decides what is sent, (C) one loop hands the requests to the devices. */
  PRB_HD void
  execute (int x, const PCand &c)
  {
    const uint32_t now = c.ts;
    const int64_t now_abs = base + now;
    const int kind = c.kind;
    int l = c.idx;
    /* (A) */
    PRec r;
    r.w0 = 0; r.uid = 0; r.aux = 0; r.s = 0; r.xw = 0;
    uint16_t ri = 0;
    if (kind == PK_RCV || kind == PK_TXC)
      {
        ri = kind == PK_RCV ? d.head[l] : d.mid[l];
        r = d.rings[ps.ring_off[l] + (ri & ps.ring_mask[l])];
        if (kind == PK_RCV) d.head[l] = (uint16_t) (ri + 1);
      }
    Pkt p;
    to_pkt (r, p);
    /* (B) */
    PCtx cur;
    Pkt s0 = p, s1 = p;
    int l0 = 0, l1 = 0, ns = 0, mode = 0;
    if (kind == PK_RCV)
      {
        /* PointToPointNetDevice::Receive (+ agent hook, ExecuteActions, IP forward, statistics) */
        cur = rcv_ctx (l, now_abs, r, ri);
        bool access = l >= st.L;
        int type = p_type (p);
        acc.ev_real++;
        if (PRB_UNLIKELY (type == PRB_DATA && p_rejected (p) && x < 23)) return;
        log_event (now_abs, access ? PRB_EV_ARR : PRB_EV_RCV, x, l, p, -1);
        if (!access)
          {
            acc.packet_hops++;
            if (type == PRB_DATA) acc.data_hops++;
          }
        int m = st.map_overlay[x];
        int data = type == PRB_DATA;
        if (m >= 0)
          {
            bool mine = p_next_hop (p) == x;
            if (type == PRB_PING_FWD)
              {
                /* PingForwardPacketManager::sendPingBackPacket */
                float delay = d2f_ (dsub_ (ns_to_seconds_ (now_abs), dmul_ ((double) p.aux, 0.001)));
                reply (s0, PRB_PING_BACK, x, p_last_hop (p), SZC_PING);
                set_p_tunnel (s0, p_tunnel (p));
                s0.uid = p.uid;
#if defined(__CUDA_ARCH__)
                s0.aux = __float_as_uint (delay);
#else
                memcpy (&s0.aux, &delay, 4);
#endif
                l0 = l ^ 1; ns = 1;
              }
            else if (data ? (mine && p_valable (p)) : (type != PRB_PING_BACK && mine && p_dst (p) == x && (type == PRB_SMALL_SIGN || p_src (p) != x)))
              {
                /* a valid Notify(): one Gym step of overlay node m */
                acc.env_steps++;
                if (data)
                  {
                    int done = p_dst (p) == x;
                    if (st.train && p_src (p) != x)
                      {
                        /* DataPacketManager::sendSmallSignalingPacket */
                        reply (s0, PRB_SMALL_SIGN, x, p_last_hop (p), st.agent_sign_szc[m]);
                        s0.uid = p.uid;
                        l0 = l ^ 1; ns = 1;
                      }
                    if (!done)
                      {
                        acc.decisions++;
                        int32_t action = policy_action (m, p);
                        if (action >= 0 && action < (int) st.agent_k[m])
                          {
                            Pkt q = p;
                            int nb = st.agent_nbr[m * st.K + action];
                            set_p_last_hop (q, x); set_p_next_hop (q, nb); set_p_ipdst (q, nb);
                            set_p_first_hop (q, 0);
                            set_p_szc (q, SZC_DATA_FWD);
                            set_p_tracked (q, 1); set_p_prev_node (q, x);
                            int lq = st.route[x * st.N + nb];
                            if (ns == 0) { s0 = q; l0 = lq; } else { s1 = q; l1 = lq; }
                            ns++;
                          }
                        else
                          acc.vanished++;
                      }
                  }
              }
          }
        /* the IP forward inside m_rxCallback and the statistics after it (point-to-point-net-device.cc:432-463) */
        if (p_first_hop (p))
          {
          }
        else if (p_ipdst (p) == x)
          {
            if (data && p_last_hop (p) == NODE_NONE && !p_valable (p) && !(p_rejected (p) && st.N + x < 23))
              acc.underlay_injected++;
          }
        else
          {
            int lq = st.route[x * st.N + p_ipdst (p)];
            if (ns == 0) { s0 = p; l0 = lq; } else { s1 = p; l1 = lq; }
            ns++;
          }
        if (data && p_dst (p) == x)
          {
            float dly = d2f_ (dsub_ (ns_to_seconds_ (now_abs), (double) p.aux));
            if (p_valable (p))
              {
                if (p_next_hop (p) == p_dst (p))
                  {
                    acc.overlay_arrived++;
                    stat_push (x, now, cur.sched, PS_OVERLAY_ARRIVED, dly);
                  }
              }
            else
              {
                acc.underlay_arrived++;
                stat_push (x, now, cur.sched, PS_UNDERLAY_ARRIVED, dly);
              }
          }
        if (type > 0 && p_dst (p) == x) acc.bytes_signaling += st.szc_bytes[p_szc (p)] - 2;
        if (data && p_last_hop (p) == NODE_NONE)
          {
            if (p_valable (p))
              {
                acc.overlay_injected++;
                acc.bytes_data += st.szc_bytes[p_szc (p)] - 2;
              }
            else
              acc.underlay_injected++;
          }
      }
    else if (kind == PK_TXC)
      {
        /* PointToPointNetDevice::TransmitComplete with a packet waiting */
        cur = txc_ctx (l, now_abs);
        l0 = l; ns = 1; mode = 1;
      }
    else if (kind == PK_FLOW)
      {
        /* PoissonGeneratorApplication::SendPacket */
        const int f = c.idx;
        acc.ev_real++;
        cur = flow_ctx (f, P_SCHED_LAZY);
        int src = st.flow_src[f], dst = st.flow_dst[f];
        blank (s0);
        set_p_type (s0, PRB_DATA);
        set_p_dst (s0, dst); set_p_src (s0, src); set_p_next_hop (s0, src); set_p_last_hop (s0, NODE_NONE);
        set_p_prev_node (s0, NODE_NONE);
        set_p_szc (s0, SZC_DATA_INJ);
        /* tag.SetStartTime (uint64_t (Now ().GetSeconds ())) (poisson-application.cc:310): the truncated double
         * equals the integer quotient for every ns count (tests/test_core_host.py pins it) */
        s0.aux = (uint32_t) ((uint64_t) now_abs / 1000000000ull);
        if (st.inject_mode == PRB_INJECT_TRACE)
          {
            int32_t i = (int32_t) d.f_pos[f];
            s0.uid = st.tr_uid[i];
            set_p_valable (s0, st.tr_flags[i] & 1);
            set_p_rejected (s0, (st.tr_flags[i] >> 1) & 1);
          }
        else
          {
            /* packet uids of synthetic traffic: (ordinal of the packet inside its flow) * F + flow */
            s0.uid = (d.f_pos[f] - 1) * (uint32_t) st.F + (uint32_t) f;
            set_p_valable (s0, st.map_overlay[src] >= 0 && st.map_overlay[dst] >= 0 ? 1 : 0);
            if (PRB_UNLIKELY (st.opt_agent && st.opt_rejected && st.opt_rejected[src * st.N + dst] > 0.0))
              set_p_rejected (s0, draw_rejected (f, src, dst));
          }
        set_p_first_hop (s0, p_valable (s0));
        set_p_ipdst (s0, dst);
        l0 = st.L + src; ns = 1;
        schedule_flow (f, false, now_abs);
        rescan_flows (x);
      }
    else
      {
        /* DataPacketManager::sendPingPackets: one ping per tunnel, built in the send loop */
        const int m = c.idx;
        acc.ev_real++;
        cur = ping_ctx (m, now_abs);
        ns = st.agent_k[m];
      }
    /* (C) */
    for (int i = 0; i < ns; i++)
      {
        int li = i == 0 ? l0 : l1;
        Pkt q = i == 0 ? s0 : s1;
        if (kind == PK_PING)
          {
            const int m = c.idx;
            int nb = st.agent_nbr[m * st.K + i];
            reply (q, PRB_PING_FWD, x, nb, SZC_PING);
            set_p_tunnel (q, i);
            q.uid = d.ping_index[m]; q.aux = (uint32_t) (now_abs / 1000000);
            li = st.route[x * st.N + nb];
          }
        dev_send (x, li, now, q, cur, (uint32_t) i, mode);
      }
    if (kind == PK_PING)
      {
        d.ping_index[c.idx] += 1;
        d.t_ping[c.idx] = now_abs + st.ping_interval_ns;
      }
  }

  PRB_HDN int
  draw_rejected (int f, int src, int dst) const
  {
    uint64_t eid = st.env_id_base + (uint64_t) env;
    uint32_t r[4];
    philox4x32 ((uint32_t) eid, (uint32_t) (eid >> 32), (uint32_t) (src * st.N + dst), d.f_pos[f] - 1,
                (uint32_t) st.seed, (uint32_t) (st.seed >> 32), r);
    double u = dmul_ (dadd_ ((double) r[2], 0.5), 1.0 / 4294967296.0);
    return u < st.opt_rejected[src * st.N + dst] ? 1 : 0;
  }

  /* what sim.cc schedules before Simulator::Run, for the nodes of this lane */
  PRB_HDN void
  start_node (int x)
  {
    int m = st.map_overlay[x];
    if (m >= 0) d.t_ping[m] = st.ping_interval_ns;
    for (int f = ps.flow_begin[x]; f < ps.flow_begin[x + 1]; f++) schedule_flow (f, true, 0);
    rescan_flows (x);
  }
};

/* New run: the 32-bit offsets of one environment are re-based on the run's start.  The work of the lanes of
 * a team is split by link (re-base) and by node (p_rebase_node, after every link is done). */
PRB_HD void
p_rebase_link (const PStatic &ps, PDyn &d, int l, uint32_t delta)
{
  uint32_t b = d.busy[l];
  d.busy[l] = b > delta ? b - delta : 0;
  for (int k = 0; k < ps.W; k++) d.w_ts[l * ps.W + k] -= delta;       /* live slots hold arrivals >= the new base */
}

/* fold one statistics record into the float32 sums (ComputeStats push order) */
PRB_HD void
p_apply_stat (prb_counters &c, const PStat &s)
{
  switch (s.code >> 30)
    {
    case PS_OVERLAY_ARRIVED:
      c.sum_cost = fadd_ (c.sum_cost, s.v); c.n_cost++;
      c.sum_e2e = fadd_ (c.sum_e2e, s.v); c.n_e2e++;
      break;
    case PS_UNDERLAY_ARRIVED:
      c.sum_underlay_cost = fadd_ (c.sum_underlay_cost, s.v); c.n_underlay_cost++;
      c.sum_underlay_e2e = fadd_ (c.sum_underlay_e2e, s.v); c.n_underlay_e2e++;
      break;
    case PS_OVERLAY_DROP:
      c.sum_cost = fadd_ (c.sum_cost, s.v); c.n_cost++;
      break;
    default:
      c.sum_underlay_cost = fadd_ (c.sum_underlay_cost, s.v); c.n_underlay_cost++;
      break;
    }
}

/* add a lane's counters to the environment's totals (device: shared/global atomics) */
PRB_HD void
p_add_u64 (int64_t *dst, uint32_t v)
{
  if (v == 0) return;
#if defined(__CUDA_ARCH__)
  atomicAdd ((unsigned long long *) dst, (unsigned long long) v);
#else
  *dst += v;
#endif
}

PRB_HD void
p_flush_acc (PScal *sc, const PAcc &a)
{
  prb_counters &c = sc->c;
  p_add_u64 (&c.overlay_injected, a.overlay_injected); p_add_u64 (&c.overlay_arrived, a.overlay_arrived);
  p_add_u64 (&c.overlay_lost, a.overlay_lost); p_add_u64 (&c.underlay_injected, a.underlay_injected);
  p_add_u64 (&c.underlay_arrived, a.underlay_arrived); p_add_u64 (&c.underlay_lost, a.underlay_lost);
  p_add_u64 (&c.bytes_data, a.bytes_data); p_add_u64 (&c.bytes_signaling, a.bytes_signaling);
  p_add_u64 (&c.packet_hops, a.packet_hops); p_add_u64 (&c.data_hops, a.data_hops);
  p_add_u64 (&c.env_steps, a.env_steps); p_add_u64 (&c.decisions, a.decisions); p_add_u64 (&c.vanished, a.vanished);
  p_add_u64 (&sc->ev_real, a.ev_real); p_add_u64 (&sc->tx_started, a.tx_started);
  if (a.errors)
    {
#if defined(__CUDA_ARCH__)
      atomicOr ((int *) &c.errors, (int) a.errors);
#else
      c.errors |= (int32_t) a.errors;
#endif
    }
}

/* read-out: same fields as the serial engine's finalize_counters */
PRB_HD void
p_finalize_counters (const Static &st, const PDyn &d, prb_counters &c)
{
  const PScal *sc = d.sc;
  c = sc->c;
  /* events = Receive + injections + ping timers + every TransmitComplete that lies before the frontier */
  int64_t pending = 0;
  for (int l = 0; l < st.LA; l++)
    if (d.busy[l] != 0 && sc->base + (int64_t) d.busy[l] - 1 >= sc->front_ts) pending++;
  c.events = sc->ev_real + sc->tx_started - pending;
  c.end_ns = sc->front_ts;
  c.avg_e2e = c.n_e2e ? c.sum_e2e / (float) c.n_e2e : 0.0f;
  c.avg_cost = c.n_cost ? c.sum_cost / (float) c.n_cost : 0.0f;
  c.avg_underlay_e2e = c.n_underlay_e2e ? c.sum_underlay_e2e / (float) c.n_underlay_e2e : 0.0f;
  c.avg_underlay_cost = c.n_underlay_cost ? c.sum_underlay_cost / (float) c.n_underlay_cost : 0.0f;
  c.signaling_overhead = c.bytes_data ? (float) (int32_t) c.bytes_signaling / (float) (int32_t) c.bytes_data : 0.0f;
}

}  // namespace prb
