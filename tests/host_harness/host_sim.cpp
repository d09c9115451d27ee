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
                     const uint8_t *valable, const uint8_t *rejected)
{
  if (!build_trace (s->hm, n, t_ns, src, dst, uid, valable, rejected, s->tr)) return -1;
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
double hs_ns_to_seconds (hs_sim *s, int64_t ns) { return ns_to_seconds (s->hm.st, ns); }
int64_t hs_tx_time_ns (uint32_t bytes, uint64_t bps) { return host_tx_time_ns (bytes, bps); }
void hs_invert_1e9 (uint64_t *hi, uint64_t *lo) { host_invert_1e9 (hi, lo); }
int32_t hs_global_routes (int32_t n, const uint8_t *adj, int32_t *out) { return host_global_routes (n, adj, out); }
int32_t hs_sizeof_step (void) { return (int32_t) sizeof (prb_step_rec); }
int32_t hs_sizeof_event (void) { return (int32_t) sizeof (prb_event); }
int32_t hs_sizeof_counters (void) { return (int32_t) sizeof (prb_counters); }
}
