/* TEST INFRASTRUCTURE — compiles the kernel body (prisma_b200/csrc/sim_core.cuh) with g++ and runs ONE
 * environment serially on the CPU, so that the event logic can be single-stepped against the oracle in
 * the GPU-less build container (tests/test_core_host.py).  This is a debugging aid of the test-suite:
 * it is not linked into libprisma_b200.so, not importable from prisma_b200/, and nothing in the product
 * falls back to it.
 */
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../prisma_b200/csrc/model.h"

using namespace prb;

struct hs_sim
{
  HostModel hm;
  HostTrace tr;
  std::vector<int32_t> pol;
  std::vector<char> blob;
  std::vector<Pkt> pool;
  std::vector<prb_event> ev;
  std::vector<prb_step_rec> steps;
  Dyn d;
  int env_index;
};

extern "C" {

hs_sim *
hs_create (const prb_config *cfg, int env_index, int64_t ev_cap, int64_t st_cap)
{
  hs_sim *s = new hs_sim ();
  if (!build_model (*cfg, s->hm))
    {
      fprintf (stderr, "hs_create: %s\n", s->hm.error.c_str ());
      delete s;
      return nullptr;
    }
  Static &st = s->hm.st;
  s->env_index = env_index;
  s->blob = s->hm.blob0;
  s->pool.resize (st.P);
  for (int i = 0; i < st.P; i++) s->pool[i].next = (uint16_t) (i + 1 < st.P ? i + 1 : NIL);
  s->ev.resize (ev_cap > 0 ? ev_cap : 1);
  s->steps.resize (st_cap > 0 ? st_cap : 1);
  st.ev_cap = ev_cap; st.st_cap = st_cap;
  bind_state (st, s->blob.data (), s->d);
  s->d.pool = s->pool.data ();
  s->d.ev_log = s->ev.data ();
  s->d.st_log = s->steps.data ();
  s->d.sc->pol_rng = 0;
  return s;
}

void hs_destroy (hs_sim *s) { delete s; }

int
hs_set_inject_trace (hs_sim *s, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst, const int64_t *uid,
                     const uint8_t *valable, const uint8_t *rejected, const uint8_t *ptype)
{
  if (!build_trace (s->hm, n, t_ns, src, dst, uid, valable, rejected, s->tr, ptype)) return -1;
  Static &st = s->hm.st;
  st.inject_mode = PRB_INJECT_TRACE;
  st.tr_n = n; st.tr_t = s->tr.t.data (); st.tr_uid = s->tr.uid.data (); st.tr_flags = s->tr.flags.data ();
  st.tr_next = s->tr.next.data (); st.flow_first = s->hm.flow_first.data ();
  return 0;
}

int
hs_set_policy (hs_sim *s, int kind, const int32_t *data, int64_t n, uint64_t seed)
{
  Static &st = s->hm.st;
  s->pol.assign (data, data + (data ? n : 0));
  st.policy = kind; st.pol_data = s->pol.data (); st.pol_n = n; st.policy_seed = seed;
  uint64_t r = (seed + (uint64_t) s->env_index) * 0x9E3779B97F4A7C15ULL + 0x1234567ULL;
  s->d.sc->pol_rng = r ? r : 1;
  s->d.sc->pol_pos = 0;
  return 0;
}

void
hs_run (hs_sim *s, int64_t until_ns)
{
  s->hm.st.mode = MODE_RUN;
  run_env_serial<true, false> (s->hm.st, s->d, s->env_index, until_ns, 0, nullptr);
}

/* Gym mode: one Notify() per call; `action` answers the previous record */
void
hs_step (hs_sim *s, int32_t action, prb_step_rec *out)
{
  s->hm.st.mode = MODE_GYM;
  run_env_serial<true, true> (s->hm.st, s->d, s->env_index, s->hm.st.sim_time_ns, action, out);
}

/* the record-free fast path of fused runs (no log): counters only */
void
hs_run_nolog (hs_sim *s, int64_t until_ns)
{
  s->hm.st.mode = MODE_RUN;
  run_env_serial<false, false> (s->hm.st, s->d, s->env_index, until_ns, 0, nullptr);
}

void
hs_counters (hs_sim *s, prb_counters *out)
{
  finalize_counters (s->hm.st, s->d, *out);
}

int64_t hs_num_events (hs_sim *s) { return s->d.sc->ev_n; }
const prb_event *hs_events (hs_sim *s) { return s->ev.data (); }
int64_t hs_num_steps (hs_sim *s) { return s->d.sc->st_n; }
const prb_step_rec *hs_steps (hs_sim *s) { return s->steps.data (); }
int32_t hs_num_links (hs_sim *s) { return s->hm.st.L; }
int32_t hs_queue_bytes (hs_sim *s, int32_t l) { return (int32_t) s->d.q_bytes[l]; }
int64_t hs_seconds_to_ns (double v) { return host_seconds_to_ns (v); }
double hs_ns_to_seconds (hs_sim *s, int64_t ns) { (void) s; return ns_to_seconds_ (ns); }
int64_t hs_tx_time_ns (uint32_t bytes, uint64_t bps) { return host_tx_time_ns (bytes, bps); }
void hs_invert_1e9 (uint64_t *hi, uint64_t *lo) { host_invert_1e9 (hi, lo); }
int32_t hs_global_routes (int32_t n, const uint8_t *adj, int32_t *out) { return host_global_routes (n, adj, out); }
int32_t hs_sizeof_step (void) { return (int32_t) sizeof (prb_step_rec); }
int32_t hs_sizeof_event (void) { return (int32_t) sizeof (prb_event); }
int32_t hs_sizeof_counters (void) { return (int32_t) sizeof (prb_counters); }
}

/* ------------------------------------------------------------------------------------------------------
 * Node-parallel engine (prisma_b200/csrc/par_core.cuh): the lanes of one team emulated one after the
 * other.  Phase 2 of a round (the handlers) runs the lanes in a SHUFFLED order: the result must not
 * depend on it, which is what makes the concurrent execution on the device legal.
 * ---------------------------------------------------------------------------------------------------- */
struct hp_sim
{
  HostModel hm;
  PModel pm;
  HostTrace tr;
  std::vector<int32_t> pol, pol_off;
  std::vector<char> blob;
  std::vector<PRec> rings;
  std::vector<prb_event> ev;
  PDyn d;
  int env_index, T;
  uint64_t rng;
  int64_t rounds, executed, chunk_ns;
};

template <bool LOG>
static void
hp_run_chunk (hp_sim *s, int64_t until_ns)
{
  const Static &st = s->hm.st;
  const PStatic &ps = s->pm.ps;
  PDyn &d = s->d;
  PScal *sc = d.sc;
  const int N = st.N, T = s->T;
  if (sc->finished) return;
  std::vector<PAcc> accs (T);
  memset (accs.data (), 0, sizeof (PAcc) * T);
  /* a new run re-bases the 32-bit offsets on its start */
  if (sc->base != sc->front_ts)
    {
      uint32_t delta = (uint32_t) (sc->front_ts - sc->base);
      for (int l = 0; l < st.LA; l++) p_rebase_link (ps, d, l, delta);
      sc->base = sc->front_ts;
      for (int x = 0; x < N; x++)
        {
          PSim<LOG> sim (st, ps, d, accs[x % T], s->env_index);
          sim.set_window (until_ns);
          sim.rescan_flows (x);
        }
    }
  if (!sc->started)
    {
      for (int x = 0; x < N; x++)
        {
          PSim<LOG> sim (st, ps, d, accs[x % T], s->env_index);
          sim.set_window (until_ns);
          sim.start_node (x);
        }
      sc->started = 1;
    }
  std::vector<PCand> cand (N);
  std::vector<int> chosen (T), order (T);
  auto fold = [&] (uint32_t bound) {
    for (;;)
      {
        int bx = -1;
        uint32_t bt = P_FAR32, bd = 0;
        /* every lane offers the oldest record of its nodes (first node wins inside a lane); same timestamp
         * on several lanes: the one that was scheduled earliest, then the lowest lane */
        for (int g = 0; g < T; g++)
          {
            int lx = -1;
            uint32_t lt = P_FAR32;
            for (int x = g; x < N; x += T)
              if (d.s_head[x] != d.s_tail[x])
                {
                  uint32_t ts = d.sring[x * P_SR + (d.s_head[x] & (P_SR - 1))].ts;
                  if (ts < bound && ts < lt)
                    {
                      lt = ts; lx = x;
                    }
                }
            if (lx < 0) continue;
            uint32_t dd = d.sring[lx * P_SR + (d.s_head[lx] & (P_SR - 1))].code & 0x3FFFFFFFu;
            if (bx < 0 || lt < bt || (lt == bt && dd > bd))
              {
                bt = lt; bx = lx; bd = dd;
              }
          }
        if (bx < 0) break;
        p_apply_stat (sc->c, d.sring[bx * P_SR + (d.s_head[bx] & (P_SR - 1))]);
        d.s_head[bx]++;
      }
  };
  for (;;)
    {
      bool any = false;
      uint32_t gmin = P_FAR32;
      for (int x = 0; x < N; x++)
        {
          PSim<LOG> sim (st, ps, d, accs[x % T], s->env_index);
          sim.set_window (until_ns);
          cand[x] = sim.select (x);
          d.t_next[x] = cand[x].ts;
          gmin = std::min (gmin, cand[x].ts);
          any = any || sim.in_time (cand[x]);
        }
      fold (any ? gmin : P_FAR32);
      if (!any) break;
      int nsafe = 0;
      for (int g = 0; g < T; g++)
        {
          chosen[g] = -1;
          for (int x = g; x < N; x += T)
            {
              PSim<LOG> sim (st, ps, d, accs[g], s->env_index);
              sim.set_window (until_ns);
              if (sim.safe (x, cand[x], gmin) && (chosen[g] < 0 || cand[x].ts < cand[chosen[g]].ts)) chosen[g] = x;
            }
          if (chosen[g] >= 0) nsafe++;
          order[g] = g;
        }
      if (nsafe == 0)
        {
          sc->c.errors |= PRB_ERR_SEQ;
          break;
        }
      for (int g = T - 1; g > 0; g--)
        {
          s->rng ^= s->rng >> 12; s->rng ^= s->rng << 25; s->rng ^= s->rng >> 27;
          int j = (int) ((s->rng * 2685821657736338717ULL) % (uint64_t) (g + 1));
          std::swap (order[g], order[j]);
        }
      for (int i = 0; i < T; i++)
        {
          int g = order[i];
          if (chosen[g] < 0) continue;
          PSim<LOG> sim (st, ps, d, accs[g], s->env_index);
          sim.set_window (until_ns);
          sim.execute (chosen[g], cand[chosen[g]]);
          s->executed++;
        }
      s->rounds++;
    }
  for (int g = 0; g < T; g++) p_flush_acc (sc, accs[g]);
  const int64_t stop = st.sim_time_ns;
  sc->front_ts = std::min (until_ns, stop);
  if (until_ns >= stop) sc->finished = 1;
}

/* like prb_run: launches cover at most max_chunk_ns each */
template <bool LOG>
static void
hp_run_t (hp_sim *s, int64_t until_ns)
{
  const int64_t stop = s->hm.st.sim_time_ns;
  const int64_t target = std::min (until_ns, stop);
  int64_t chunk = s->chunk_ns > 0 ? std::min (s->chunk_ns, s->pm.ps.max_chunk_ns) : s->pm.ps.max_chunk_ns;
  do
    {
      int64_t front = s->d.sc->front_ts;
      int64_t end = std::min (target, front + chunk);
      hp_run_chunk<LOG> (s, end >= stop ? std::max (until_ns, stop) : end);
    }
  while (s->d.sc->front_ts < target && !s->d.sc->finished);
}

extern "C" {

hp_sim *
hp_create (const prb_config *cfg, int env_index, int64_t ev_cap, int team)
{
  hp_sim *s = new hp_sim ();
  if (!build_model (*cfg, s->hm) || !build_pmodel (*cfg, s->hm, s->pm))
    {
      fprintf (stderr, "hp_create: %s %s\n", s->hm.error.c_str (), s->pm.why.c_str ());
      delete s;
      return nullptr;
    }
  Static &st = s->hm.st;
  s->env_index = env_index;
  s->T = team;
  s->rng = 0x9E3779B97F4A7C15ULL + (uint64_t) env_index;
  s->rounds = s->executed = 0;
  s->chunk_ns = 0;
  s->blob = s->pm.pblob0;
  s->rings.resize (s->pm.ps.ring_records);
  s->ev.resize (ev_cap > 0 ? ev_cap : 1);
  st.ev_cap = ev_cap; st.st_cap = 0;
  p_bind_state (s->pm.ps, s->blob.data (), s->d);
  s->d.rings = s->rings.data ();
  s->d.ev_log = s->ev.data ();
  return s;
}

void hp_destroy (hp_sim *s) { delete s; }

int
hp_set_inject_trace (hp_sim *s, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst, const int64_t *uid,
                     const uint8_t *valable, const uint8_t *rejected)
{
  if (!build_trace (s->hm, n, t_ns, src, dst, uid, valable, rejected, s->tr)) return -1;
  Static &st = s->hm.st;
  st.inject_mode = PRB_INJECT_TRACE;
  st.tr_n = n; st.tr_t = s->tr.t.data (); st.tr_uid = s->tr.uid.data (); st.tr_flags = s->tr.flags.data ();
  st.tr_next = s->tr.next.data (); st.flow_first = s->hm.flow_first.data ();
  s->pm.ps.tr_prev = s->tr.prev.data ();
  return 0;
}

/* kind TABLE: data[M*M]; kind NODE_TRACE: data = [M+1 offsets][actions of node 0][actions of node 1]... */
int
hp_set_policy (hp_sim *s, int kind, const int32_t *data, int64_t n)
{
  Static &st = s->hm.st;
  if (kind == PRB_POLICY_NODE_TRACE)
    {
      s->pol_off.assign (data, data + st.M + 1);
      s->pol.assign (data + st.M + 1, data + n);
      s->pm.ps.pol_off = s->pol_off.data ();
    }
  else
    s->pol.assign (data, data + (data ? n : 0));
  st.policy = kind; st.pol_data = s->pol.data (); st.pol_n = (int64_t) s->pol.size ();
  return 0;
}

void
hp_run (hp_sim *s, int64_t until_ns, int log)
{
  if (log) hp_run_t<true> (s, until_ns); else hp_run_t<false> (s, until_ns);
}

void hp_counters (hp_sim *s, prb_counters *out) { p_finalize_counters (s->hm.st, s->d, *out); }
int64_t hp_num_events (hp_sim *s) { return std::min<int64_t> (s->d.sc->ev_n, (int64_t) s->ev.size ()); }
const prb_event *hp_events (hp_sim *s) { return s->ev.data (); }
int32_t hp_num_links (hp_sim *s) { return s->hm.st.L; }
int32_t hp_queue_bytes (hp_sim *s, int32_t l) { return (int32_t) s->d.q_bytes[l]; }
int64_t hp_rounds (hp_sim *s) { return s->rounds; }
int64_t hp_executed (hp_sim *s) { return s->executed; }
int32_t hp_wire_slots (hp_sim *s) { return s->pm.ps.W; }
int32_t hp_state_bytes (hp_sim *s) { return s->pm.ps.pblob_bytes; }
int64_t hp_lookahead_ns (hp_sim *s) { return s->pm.ps.la_ns; }
void hp_set_chunk (hp_sim *s, int64_t ns) { s->chunk_ns = ns; }
}
