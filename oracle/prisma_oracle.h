/* TEST INFRASTRUCTURE — CPU restatement of the PRISMA / ns-3 environment step loop.
 *
 * This is the parity oracle of prisma_b200: a plain-C, single-environment,
 * strictly sequential discrete-event simulator that restates what the
 * reference's ns-3 program does on the hot path.  It is NOT
 * product code: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call it.
 *
 * Parity status: PINNED.  The restatement is checked event-by-event against
 * the reference's own C++ (oracle/_ref/prisma_ref, built by oracle/build_ref.py
 * from the sources under /root/reference) — see tests/test_oracle_vs_ref.py
 * and the committed fixtures under tests/golden/.
 *
 * Every function cites the reference file:line it follows
 * (paths relative to /root/reference).
 */
#ifndef PRISMA_ORACLE_H
#define PRISMA_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PO_MAX_NODES 128
#define PO_MAX_K 16
#define PO_LASTHOP_NONE 1000 /* poisson-application.cc:309 */

/* packet types: prisma/ns3/enum-and-constants.h:5-11 */
enum { PO_DATA = 0, PO_BIG_SIGN = 1, PO_SMALL_SIGN = 2, PO_PING_FWD = 3, PO_PING_BACK = 4 };

/* event-log record kinds (same vocabulary as oracle/ref_harness/ref_harness.cc) */
enum { PO_EV_INJ = 0, PO_EV_ARR = 1, PO_EV_ENQ = 2, PO_EV_DRP = 3, PO_EV_TXB = 4, PO_EV_RCV = 5 };

enum { PO_POLICY_ZERO = 0, PO_POLICY_TABLE = 1, PO_POLICY_TRACE = 2, PO_POLICY_RANDOM = 3 };
enum { PO_INJECT_TRACE = 0, PO_INJECT_PHILOX = 1 };

typedef struct po_config
{
  int32_t n_nodes;                 /* N underlay nodes                       sim.cc:303-305 */
  int32_t n_overlay;               /* M overlay nodes                        sim.cc:304      */
  const uint8_t *phys_adj;         /* [N*N] 0/1                              sim.cc:274-276  */
  const uint8_t *overlay_adj;      /* [M*M] 0/1                              sim.cc:277-279  */
  const int32_t *map_overlay;      /* [N] overlay index or -1                sim.cc:281-282  */
  const uint64_t *traffic_bps;     /* [N*N] ceil(rate*load_factor), 0 = no flow   sim.cc:623 */
  const int32_t *route_out;        /* [N*N] out directed-link id at node x towards 10.2.2.(d+1);
                                      NULL = compute with po_global_routes        sim.cc:683 */
  const double *opt_rejected;      /* [N*N] or NULL                          sim.cc:284-285  */
  uint64_t link_rate_bps;          /* LinkRate                               sim.cc:422      */
  int64_t link_delay_ns;           /* LinkDelay                              sim.cc:423      */
  uint32_t packet_size;            /* AvgPacketSize (payload bytes)          sim.cc:150      */
  uint32_t max_out_buffer_size;    /* MaxBufferLength in bytes               sim.cc:424      */
  int64_t sim_time_ns;             /* Simulator::Stop                        sim.cc:703      */
  int64_t ping_interval_ns;        /* Seconds(float pingPacketIntervalTime)  data-packet-manager.cc:117-122 */
  uint32_t moving_average_size;    /* movingAverageObsSize                   sim.cc:549      */
  int32_t ping_as_obs;             /* pingAsObs                              sim.cc:548      */
  int32_t activate_underlay_traffic; /*                                      sim.cc:505-507  */
  int32_t train;                   /* small signalling per hop               packet-routing-gym.cc:205-208 */
  int32_t signaling_type;          /* 0 ideal, 1 NN, 2 target, 3 digital_twin      sim.cc:371-393 */
  int32_t opt_agent;               /* AgentType=="opt"                       sim.cc:376-377  */
  double loss_penalty;             /* lossPenalty                            sim.cc:246-247  */
  int32_t inject_mode;             /* PO_INJECT_*                                            */
  uint64_t seed;                   /* PO_INJECT_PHILOX: stream key                           */
  uint64_t env_id;                 /* PO_INJECT_PHILOX: which environment of the batch       */
  /* NN-weight segments between overlay neighbours (train && signaling && signalingType=="NN"):
   * sim.cc:634-647, big-signaling-application.cc:223-310.  All zero = no such sources. */
  int32_t signaling;               /* activateSignaling after sim.cc:374-376                 */
  uint32_t big_signaling_size;     /* bigSignalingSize, bytes                sim.cc:144      */
  double sync_step;                /* syncStep, seconds                      sim.cc:143      */
} po_config;

typedef struct po_event
{
  int64_t t_ns;
  int32_t kind;      /* PO_EV_* */
  int32_t node;      /* node at which the record is taken (traffic node of i = N+i) */
  int32_t link;      /* directed link id (access link of node i = L+i) */
  int64_t uid;       /* data packets: packet uid; others: -1 */
  int32_t size;      /* bytes incl. PPP */
  int32_t type, src, dst, last_hop, next_hop, valable;
  int32_t qbytes;    /* ENQ/DRP/TXB: queue bytes after the operation, else -1 */
} po_event;

typedef struct po_decision
{
  int64_t t_ns;
  int32_t node;       /* underlay id of the overlay node stepping */
  int32_t overlay;    /* overlay index (the Gym interface index)  */
  int64_t uid;        /* data: packet uid; small signalling: uid it reports; big signalling: NNIndex << 32 | segIndex */
  int32_t ptype;
  int32_t done;
  int32_t action;     /* -1 when the Forwarder ignores the policy (forwarder.py:149-150) */
  int32_t nobs;
  int32_t obs[1 + PO_MAX_K];
  /* restated Forwarder bookkeeping (forwarder.py:334-379) */
  int32_t is_new;     /* first time the agents see this packet */
  int32_t prev_node;  /* underlay id of the previous decision maker, -1 if new; big signalling: the signalling node */
  double hop_time;    /* now - time of previous decision, seconds (exact ns difference) */
  float e2e_delay;    /* info field 0: now_s - start_s (packet-manager.cc:121-122) */
  int32_t n_lost;     /* lost uids flushed with this step (data-packet-manager.cc:229-237) */
  int64_t lost_uid[8];
} po_decision;

typedef struct po_counters
{
  int64_t overlay_injected, overlay_arrived, overlay_lost;
  int64_t underlay_injected, underlay_arrived, underlay_lost;
  int64_t bytes_data, bytes_signaling;
  int64_t n_e2e, n_cost, n_underlay_e2e, n_underlay_cost;
  float sum_e2e, sum_cost, sum_underlay_e2e, sum_underlay_cost;   /* float32 running sums in push order (compute-stats-v2.cc:87-97) */
  float avg_e2e, avg_cost, avg_underlay_e2e, avg_underlay_cost, signaling_overhead;
  int64_t packet_hops;      /* Receive() calls on switch<->switch devices, any type */
  int64_t data_hops;        /* same, data packets only */
  int64_t env_steps;        /* valid Notify() calls */
  int64_t decisions;        /* steps where the policy was consulted */
  int64_t events;           /* simulator events executed */
  int64_t vanished;         /* action >= k (data-packet-manager.cc:294-297) */
  int64_t end_ns;
} po_counters;

typedef struct po_sim po_sim;

/* ---- ns-3 arithmetic (core/model/nstime.h, int64x64-128.{h,cc}, network/utils/data-rate.cc) ---- */
int64_t po_seconds_to_ns (double s);
double po_ns_to_seconds (int64_t ns);
int64_t po_tx_time_ns (uint32_t bytes, uint64_t bps);
void po_invert_1e9 (uint64_t *hi, uint64_t *lo);

/* ---- topology helpers ---- */
int32_t po_num_links (int32_t n, const uint8_t *adj);
/* directed link table in device-creation order (sim.cc:414-433): link 2e = i->j, 2e+1 = j->i */
int32_t po_build_links (int32_t n, const uint8_t *adj, int32_t *from, int32_t *to);
/* restatement of GlobalRouteManagerImpl SPF + Ipv4GlobalRouting::LookupGlobal first-route choice */
int32_t po_global_routes (int32_t n, const uint8_t *adj, int32_t *route_out);

/* ---- simulator ---- */
po_sim *po_create (const po_config *cfg);
void po_destroy (po_sim *s);
void po_set_event_log (po_sim *s, int32_t enable);
int32_t po_set_inject_trace (po_sim *s, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst,
                             const int64_t *uid, const uint8_t *valable, const uint8_t *rejected);
/* the same with the packet type of every injection (PO_DATA / PO_BIG_SIGN); ptype == NULL: all data */
int32_t po_set_inject_trace_typed (po_sim *s, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst,
                                   const int64_t *uid, const uint8_t *valable, const uint8_t *rejected,
                                   const uint8_t *ptype);
int32_t po_set_policy (po_sim *s, int32_t kind, const int32_t *data, int64_t n, uint64_t seed);
int32_t po_run (po_sim *s, int64_t until_ns);
void po_get_counters (po_sim *s, po_counters *out);
int64_t po_num_events (po_sim *s);
const po_event *po_events (po_sim *s);
int64_t po_num_decisions (po_sim *s);
const po_decision *po_decisions (po_sim *s);
int32_t po_queue_bytes (po_sim *s, int32_t link);
int32_t po_num_links_of (po_sim *s);
const int32_t *po_route_table (po_sim *s);

/* Philox-4x32-10 (Salmon et al. 2011), counter-based generator shared by the synthetic injection spec */
void po_philox4x32 (const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif
