/* prisma_b200 — C ABI of the batched, device-resident packet-routing environment (libprisma_b200.so).
 *
 * This is the drop-in boundary for PRISMA's environment step loop.  What each entry point replaces
 * in the reference (paths relative to the reference checkout):
 *
 *   prb_create / prb_destroy   the ns-3 process launch + topology build
 *                              (prisma/source/run_ns3.py:7-78, prisma/ns3/sim.cc:122-700) and the
 *                              per-node ZMQ bind (prisma/ns3_model/ns3env.py:41-60,397-402)
 *   prb_reset                  Ns3Env.reset (prisma/ns3_model/ns3env.py:422-426) for every env
 *   prb_step                   Ns3Env.step (ns3env.py:417-420) == Ns3ZmqBridge.send_actions +
 *                              rx_env_state (ns3env.py:179-247) == OpenGymInterface::NotifyCurrentState
 *                              (ns3-gym/src/opengym/model/opengym_interface.cc:213-286): one call advances
 *                              EVERY environment to its next valid Notify()
 *   prb_run                    the same loop with the agent hook evaluated on the device
 *                              (Forwarder.take_action, prisma/source/forwarder.py:171-195): sp / table agents
 *   prb_set_policy             Forwarder agent selection (forwarder.py:171-195, argument_parser.py:64-70)
 *   prb_set_inject_trace       replaces PoissonGeneratorApplication's RNG draws by a recorded trace
 *                              (prisma/ns3/poisson-application.cc:265-358) — parity / replay mode
 *   prb_counters               PacketManager::getInfo's global statistics
 *                              (prisma/ns3/packet-manager.cc:119-176, compute-stats-v2.cc:115-266)
 *   prb_seconds_to_ns ...      ns-3 Time / DataRate arithmetic used to derive the integer-ns inputs
 *                              (ns3-gym/src/core/model/nstime.h:235-236,456-475, int64x64-128.{h,cc},
 *                               ns3-gym/src/network/utils/data-rate.cc:235-240)
 *
 * Conventions: every call returns 0 or a negative prb_status; prb_last_error() gives the message of the
 * last failure on the calling thread.  A handle is bound to one CUDA device and one stream; handles are
 * not thread-safe, distinct handles are independent.  Pointers are HOST pointers unless the parameter
 * name ends in _dev.  There is no CPU execution path: without a CUDA device prb_create fails.
 */
#ifndef PRISMA_B200_H
#define PRISMA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PRB_MAX_K 16          /* overlay degree limit (observation = 1 + k words) */
#define PRB_MAX_NODES 250
#define PRB_LASTHOP_NONE 1000 /* prisma/ns3/poisson-application.cc:309 */
#define PRB_MAX_LOST 8

enum prb_status { PRB_OK = 0, PRB_EINVAL = -1, PRB_ECUDA = -2, PRB_ENOMEM = -3, PRB_ESTATE = -4, PRB_EOVERFLOW = -5 };

/* packet types: prisma/ns3/enum-and-constants.h:5-11 */
enum { PRB_DATA = 0, PRB_BIG_SIGN = 1, PRB_SMALL_SIGN = 2, PRB_PING_FWD = 3, PRB_PING_BACK = 4 };
/* event-log record kinds */
enum { PRB_EV_INJ = 0, PRB_EV_ARR = 1, PRB_EV_ENQ = 2, PRB_EV_DRP = 3, PRB_EV_TXB = 4, PRB_EV_RCV = 5 };
/* agent hooks evaluated on the device (prb_run) */
enum { PRB_POLICY_ZERO = 0, PRB_POLICY_TABLE = 1, PRB_POLICY_TRACE = 2, PRB_POLICY_RANDOM = 3, PRB_POLICY_NODE_TRACE = 4 };
/* which engine advances the environments in prb_run (prb_set_engine) */
enum { PRB_ENGINE_AUTO = 0, PRB_ENGINE_SERIAL = 1, PRB_ENGINE_PARALLEL = 2 };
enum { PRB_INJECT_TRACE = 0, PRB_INJECT_PHILOX = 1 };
/* env error bits (prb_errors) */
enum { PRB_ERR_POOL = 1, PRB_ERR_LOG = 2, PRB_ERR_PING_WINDOW = 4, PRB_ERR_TRACE = 8, PRB_ERR_SEQ = 16, PRB_ERR_TIE = 32 };

typedef struct prb_config
{
  int32_t num_envs;                /* E independent environments on this handle                         */
  int32_t n_nodes;                 /* N underlay nodes                                   sim.cc:303-305 */
  int32_t n_overlay;               /* M overlay nodes                                    sim.cc:304     */
  const uint8_t *phys_adj;         /* [N*N] physical adjacency                           sim.cc:274-276 */
  const uint8_t *overlay_adj;      /* [M*M] overlay adjacency                            sim.cc:277-279 */
  const int32_t *map_overlay;      /* [N] overlay index or -1                            sim.cc:281-282 */
  const uint64_t *traffic_bps;     /* [N*N] ceil(rate*load_factor), 0 = no flow          sim.cc:623     */
  const int32_t *route_out;        /* [N*N] optional: out link of x towards d, else global routing  sim.cc:683 */
  const double *opt_rejected;      /* [N*N] optional                                     sim.cc:284-285 */
  uint64_t link_cap_bps;           /* link_cap                                           sim.cc:422     */
  int64_t link_delay_ns;           /* link_delay                                         sim.cc:423     */
  uint32_t packet_size;            /* packet_size (payload bytes)                        sim.cc:150     */
  uint32_t max_out_buffer_size;    /* max_out_buffer_size, bytes                         sim.cc:424     */
  int64_t sim_time_ns;             /* simTime                                            sim.cc:703     */
  int64_t ping_interval_ns;        /* pingPacketIntervalTime      data-packet-manager.cc:117-122        */
  uint32_t moving_average_size;    /* movingAverageObsSize                               sim.cc:549     */
  int32_t ping_as_obs;             /* pingAsObs                                          sim.cc:548     */
  int32_t activate_underlay_traffic; /* activateUnderlayTraffic                          sim.cc:505-507 */
  int32_t train;                   /* train: one small-signalling packet per hop   packet-routing-gym.cc:205-208 */
  int32_t signaling_type;          /* 0 ideal, 1 NN, 2 target, 3 digital_twin            sim.cc:371-393 */
  int32_t opt_agent;               /* agent_type == "opt"                                sim.cc:376-377 */
  double loss_penalty;             /* loss_penalty                                       sim.cc:246-247 */
  uint64_t seed;                   /* synthetic traffic: Philox key                                      */
  uint64_t env_id_base;            /* synthetic traffic: id of env 0 (env e draws stream env_id_base+e)  */
  int32_t pool_capacity;           /* packet records per env, 0 = derive from the topology              */
  int32_t lanes_per_env;           /* 0 = default (32: one warp per environment)                        */
  /* NN-weight segments between overlay neighbours (BigSignalingGeneratorApplication): one source per directed
   * overlay edge that carries traffic when train && signaling && signaling_type == NN; all zero = none */
  int32_t signaling;               /* activateSignaling after sim.cc:374-376 (signalingSim, not sp/opt/ideal) */
  uint32_t big_signaling_size;     /* bigSignalingSize, bytes                            sim.cc:144,641 */
  double sync_step;                /* syncStep, seconds                                  sim.cc:143,641 */
} prb_config;

/* one record of the parity event log (same vocabulary as the reference harness) — 64 bytes */
typedef struct prb_event
{
  int64_t t_ns;
  int32_t kind;      /* PRB_EV_* */
  int32_t node;      /* node where the record is taken (traffic node of i = N+i) */
  int32_t link;      /* directed link id (access link of node i = L+i) */
  int32_t pad_;
  int64_t uid;       /* data packets: uid, others -1 */
  int32_t size;      /* bytes incl. PPP */
  int32_t type, src, dst, last_hop, next_hop, valable;
  int32_t qbytes;    /* ENQ/DRP/TXB: queue bytes after the operation, else -1 */
} prb_event;

/* one Gym step (one valid Notify()) of one environment — 208 bytes, all fields 4- or 8-byte */
typedef struct prb_step_rec
{
  int64_t t_ns;         /*   0 simulated time of the step                                              */
  int64_t uid;          /*   8 data: packet uid; small signalling: uid it reports; big signalling:
                               NNIndex << 32 | segIndex (big-signaling-packet-manager.cc:104-121)        */
  double hop_time;      /*  16 now - time of the packet's previous decision, s (forwarder.py:360)       */
  int32_t node;         /*  24 underlay id of the stepping overlay node                                 */
  int32_t overlay;      /*  28 its overlay index == the reference's per-node env index                  */
  int32_t ptype;        /*  32 PRB_DATA / PRB_SMALL_SIGN / PRB_BIG_SIGN                                  */
  int32_t done;         /*  36 getGameOver (data-packet-manager.cc:224-227)                             */
  int32_t action;       /*  40 action applied (-1: Forwarder ignores the policy, forwarder.py:149-150)  */
  int32_t need_action;  /*  44 1: prb_step consumes actions_dev[e] for this record                      */
  int32_t nobs;         /*  48                                                                          */
  int32_t is_new;       /*  52 first time the agents see this packet (forwarder.py:316-317)             */
  int32_t prev_node;    /*  56 underlay id of the previous decision maker, -1 if new; big signalling: the
                               node whose weights the segment carries ("NodeId Signaled" = its overlay index) */
  int32_t n_lost;       /*  60 lost uids flushed with this step (data-packet-manager.cc:229-237)        */
  float e2e_delay;      /*  64 info field 0 (packet-manager.cc:121-122)                                 */
  int32_t valid;        /*  68 1: this env produced a step; 0: it reached simTime (obs[0] == -1 sentinel) */
  int32_t obs[1 + PRB_MAX_K]; /* 72 */
  int32_t src;          /* 140 underlay id of the packet's source (MyTag source; "Source" of DataPacketManager::getInfo
                               is its overlay index): what an `opt` agent evaluated by the caller needs next to obs[0]
                               (prisma/source/utils.py:161-199) */
  int64_t lost_uid[PRB_MAX_LOST]; /* 144 most recent first */
} prb_step_rec;

/* per-environment statistics (compute-stats-v2.cc:59-72,115-266) — 200 bytes */
typedef struct prb_counters
{
  int64_t overlay_injected, overlay_arrived, overlay_lost;
  int64_t underlay_injected, underlay_arrived, underlay_lost;
  int64_t bytes_data, bytes_signaling;
  int64_t n_e2e, n_cost, n_underlay_e2e, n_underlay_cost;
  float sum_e2e, sum_cost, sum_underlay_e2e, sum_underlay_cost;      /* float32 sums in push order */
  float avg_e2e, avg_cost, avg_underlay_e2e, avg_underlay_cost, signaling_overhead;
  int32_t errors;           /* PRB_ERR_* bits, must be 0 */
  int64_t packet_hops;      /* Receive() on switch<->switch devices, any packet type */
  int64_t data_hops;        /* same, data packets only */
  int64_t env_steps;        /* valid Notify() calls */
  int64_t decisions;        /* steps where the agent chose a next hop */
  int64_t events;           /* simulator events executed */
  int64_t vanished;         /* action >= k (data-packet-manager.cc:294-297) */
  int64_t end_ns;           /* simulated clock */
} prb_counters;

typedef struct prb_env prb_env;

const char *prb_last_error (void);
const char *prb_version (void);

/* ---- ns-3 arithmetic on the host (inputs of prb_config are integer ns / bit/s) ---- */
int64_t prb_seconds_to_ns (double seconds);
double prb_ns_to_seconds (int64_t ns);
int64_t prb_tx_time_ns (uint32_t bytes, uint64_t bps);
/* directed links in device-creation order (sim.cc:414-433); returns L, fills from/to when non-NULL */
int32_t prb_build_links (int32_t n, const uint8_t *adj, int32_t *from, int32_t *to);
/* ns-3 global routing as the reference uses it: out link of node x towards node d, -1 on the diagonal */
int32_t prb_global_routes (int32_t n, const uint8_t *adj, int32_t *route_out);

/* ---- environment ---- */
int prb_create (const prb_config *cfg, int device, void *cuda_stream, prb_env **out);
int prb_destroy (prb_env *env);
int prb_reset (prb_env *env);
/* engine of prb_run: PRB_ENGINE_AUTO picks the node-parallel engine (one lane per network node, conservative
 * lookahead inside an environment) for table / per-node-trace agents without ordered logs, the event-serial
 * one otherwise (prb_step, trace and random agents, prb_enable_log).  team: lanes per environment of the
 * parallel engine, 0 = from the node count.  prb_engine: the engine that has advanced the environments
 * since the last reset (0 none).  prb_round_stats: rounds and events of the parallel engine so far. */
int prb_set_engine (prb_env *env, int mode, int team);
/* Environments of this episode that the node-parallel engine handed over to the event-serial engine because the
 * order of two simultaneous events could not be rebuilt from local information (rare: long lock-step chains)
 * or because a link's record ring filled up (a queue of nothing but the smallest packets).
 * They are replayed from the reset state in global (timestamp, uid) order and kept up to date before every
 * read-out, so results stay exact; with an ordered log or a per-node action trace PRB_ERR_TIE is reported instead. */
int prb_handovers (prb_env *env);
/* The same count, and the device time (ms, CUDA events on the environment's stream) the serial engine has spent replaying
 * those environments in this episode.  A table agent may be replaced between prb_run calls (prb_set_policy): the replay
 * runs every stretch of simulated time with the table the parallel launches used for it.
 * Replaces nothing in the reference (its single event list needs no such repair); diagnostics for bench.py. */
int prb_handover_stats (prb_env *env, int64_t *envs, double *replay_ms);
int prb_engine (prb_env *env);
int prb_round_stats (prb_env *env, int64_t *rounds, int64_t *executed);
int prb_num_links (prb_env *env);        /* L (switch<->switch, directed) */
int prb_state_bytes (prb_env *env, int64_t *blob_bytes, int64_t *pool_bytes);

/* traffic: recorded injections shared by every env (replay), or NULL/0 to go back to Philox */
int prb_set_inject_trace (prb_env *env, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst,
                          const int64_t *uid, const uint8_t *valable, const uint8_t *rejected);
/* the same with the packet type of every injection (PRB_DATA, or PRB_BIG_SIGN for the segments of
 * big-signaling-application.cc:261-306); ptype == NULL: all data */
int prb_set_inject_trace_typed (prb_env *env, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst,
                                const int64_t *uid, const uint8_t *valable, const uint8_t *rejected,
                                const uint8_t *ptype);
/* agent hook evaluated on the device: TABLE data[M*M] (row = stepping overlay node, column = obs[0]);
 * TRACE data[n] actions in decision order; RANDOM uses seed (+ env index); NODE_TRACE data = [M+1 offsets]
 * [actions of overlay node 0][of node 1]...: every node replays its own decisions (parallel engine) */
int prb_set_policy (prb_env *env, int kind, const int32_t *data, int64_t n, uint64_t seed);

/* parity logs: per-env capacities (records); 0 disables */
int prb_enable_log (prb_env *env, int64_t event_capacity, int64_t step_capacity);
int prb_read_events (prb_env *env, int32_t e, prb_event *out, int64_t capacity, int64_t *n);
int prb_read_steps (prb_env *env, int32_t e, prb_step_rec *out, int64_t capacity, int64_t *n);

/* fused-agent mode: advance every env until simulated time until_ns (or simTime); a call whose until_ns lies at or
 * before the time already reached does nothing (both engines) */
int prb_run (prb_env *env, int64_t until_ns);
/* Gym mode: apply actions_dev[e] (int32, device; may be NULL on the first call) to the pending step of
 * every env that asked for one, then advance every env to its next valid Notify(); out_dev[e] (device,
 * E records) receives it */
int prb_step (prb_env *env, const int32_t *actions_dev, prb_step_rec *out_dev);
/* the same through host buffers (copies inside) */
int prb_step_host (prb_env *env, const int32_t *actions, prb_step_rec *out);

/* ---- dqn_buffer agents on the device (prisma/source/models.py:257-301, learner.py:142-159) ----
 * prb_qnet_load: the weights of the M per-node networks, prb_qnet_num_floats floats, node after node, per node
 *   W_dst[M][32] b_dst[32] W_buf[K][32] b_buf[32] W_1[64][64] b_1[64] W_2[64][64] b_2[64] W_3[64][K] b_3[K]
 *   (Keras kernel layout [in][out], K = largest overlay degree, unused columns ignored).
 * prb_qnet_act: argmin-Q action (uniform random neighbour with probability eps) for the pending step of every
 *   environment; rec_dev / actions_dev are the buffers of prb_step; q_out_dev [E][K] optional.
 * prb_step_qnet: n_steps x (prb_qnet_act, prb_step) without leaving the device: out_dev must hold the records
 *   of the previous prb_step (zeroed before the first call of an episode). */
int64_t prb_qnet_num_floats (prb_env *env);
int prb_qnet_load (prb_env *env, const float *weights, int64_t n_floats);
int prb_qnet_act (prb_env *env, const prb_step_rec *rec_dev, int32_t *actions_dev, float *q_out_dev, float eps, uint64_t seed);
/* How the Q-networks are evaluated (DQN_AGENT.step, prisma/source/learner.py:142-159; the reference hands the batch to
 * TensorFlow / cuBLAS).  FUSED (default): inside the simulator kernel, one warp per environment, fp32 (prb_step_qnet is one
 * launch).  WARP: the same arithmetic as its own kernel (prb_qnet_act).  MMA: the pending decisions of all environments are
 * bucketed by node and evaluated in tiles of 16 on the tensor cores (mma.sync TF32 with split operands, fp32 accumulate:
 * Q-values within 2e-5 of the fp32 path); prb_step_qnet then issues three launches per step.  Also: PRB_QNET_MODE=mma|warp. */
enum { PRB_QNET_FUSED = 0, PRB_QNET_WARP = 1, PRB_QNET_MMA = 2 };
int prb_qnet_set_mode (prb_env *env, int mode);
/* Depth of the loaded Q-networks after the concatenation of the two input branches: 2 (DQN_buffer_model, DQN_buffer_lite_model,
 * DQN_routing_model; default) or 1 (DQN_buffer_lighter_model and its _2 / _3 siblings, prisma/source/models.py:39-181: the output layer
 * reads the first hidden layer; W_2 / b_2 of the layout are ignored).  Narrower layers are zero-padded layouts (prisma_b200/qnet.py). */
int prb_qnet_set_arch (prb_env *env, int hidden_layers);
int prb_step_qnet (prb_env *env, int32_t n_steps, float eps, uint64_t seed, prb_step_rec *out_dev);

/* ---- replay buffers in HBM (prisma/source/replay_buffer.py:12-86; producers forwarder.py:135-161,214-244,352-379) ----
 * One ring of rows_per_node transitions per overlay node, shared by all environments; a row is 2*(1+K)+3 floats
 *   obs[1+K] action reward next_obs[1+K] done.
 * prb_replay_record: feed the records of the last prb_step together with the actions that answer them (call it
 *   between choosing the actions and the next prb_step; prb_step_qnet does so itself): the transit packet's
 *   transition goes to the previous decision maker's ring with reward = hop time, lost packets to the reporting
 *   node's ring with reward = loss_penalty, next_obs = [dst, 0...], done = 1.  pending_slots (power of two) is
 *   the per-environment table of decisions in flight, keyed by packet uid (Agent.temp_obs).
 * prb_replay_sample: uniform sampling with replacement (ReplayBuffer.sample) into device arrays [batch][1+K] / [batch].
 * prb_replay_counts: rows ever written per node (host), transitions dropped because their decision was not found.
 * prb_replay_read: the stored rows of one node to the host (ReplayBuffer.save). */
int prb_replay_enable (prb_env *env, int64_t rows_per_node, int32_t pending_slots);
int prb_replay_record (prb_env *env, const prb_step_rec *rec_dev, const int32_t *actions_dev);
/* `target` signalling (prisma/source/trainer.py:52-55, forwarder.py:389-431 with signalingSim = 0): the reward column
 * of a transit transition holds the TD target that the NEXT node computes with its own (current) network when the
 * packet reaches it, hop_time + gamma * min over its actions that do not lead back (learner.py:232-255); terminal and
 * lost-packet transitions keep the plain reward.  Needs the device-side agents (prb_qnet_act before
 * prb_replay_record, or prb_step_qnet). */
int prb_replay_set_target_mode (prb_env *env, int on, float gamma);
/* Prioritized variant (prisma/source/replay_buffer.py:393-534, used with `target` signalling: forwarder.py:494-505): draws
 * stay uniform (replay_buffer.py:497); every row remembers the gradient step of the network that produced its target
 * (prb_replay_set_gradient_steps: the learners' step counters, host [M]), every node the latest step it has seen per
 * action, and prb_replay_sample_weighted returns weight = p / sum (p) * len with p = (row step / latest step) ^ alpha
 * — the state the reference reaches after its next update_priorities (until then it keeps new rows at max priority).
 * Without the prioritized variant the weights are 1.  prb_replay_read_priorities: row steps and latest steps of a node. */
int prb_replay_enable_prioritized (prb_env *env, float alpha);
int prb_replay_set_gradient_steps (prb_env *env, const int32_t *steps /* host [M] */);
int prb_replay_sample_weighted (prb_env *env, int32_t node, int32_t batch, uint64_t seed, float *obs_dev, int32_t *action_dev,
                                float *reward_dev, float *next_obs_dev, float *done_dev, float *weights_dev);
int prb_replay_read_priorities (prb_env *env, int32_t node, float *gstep_host, int64_t max_rows, int32_t *latest_host /* [K] */);
int prb_replay_counts (prb_env *env, int64_t *rows_written /* host [M] */, int64_t *missed /* host, may be NULL */);
int prb_replay_read (prb_env *env, int32_t node, float *rows_host, int64_t max_rows, int64_t *n_rows);
int prb_replay_sample (prb_env *env, int32_t node, int32_t batch, uint64_t seed, float *obs_dev, int32_t *action_dev,
                       float *reward_dev, float *next_obs_dev, float *done_dev);

/* ---- the DQN learner's gradient step on the device (dqn_buffer agents) ----
 * Replaces DQN_AGENT.train / get_neighbor_target_value (prisma/source/learner.py:162-189,231-255) and Trainer.step
 * (prisma/source/trainer.py:45-99): TD targets through the neighbours' target networks (or the stored targets with `target`
 * signalling, trainer.py:52-55), Huber loss, Adam with Keras' defaults, one optimiser state per node.  The online weights are
 * the ones prb_qnet_load installed: training updates them in place, so the device-side agents act on the new weights at once.
 * prb_replay_sample_all: ReplayBuffer.sample for every node at once into [M][batch][...] device arrays (active_dev[m] = 1 when
 *   node m holds at least min_rows transitions, trainer.py:41; weights_dev may be NULL).
 * prb_learner_grad: gradients of every active node into bucket_dev [M * P floats][M activity flags] — ONE buffer, the unit of the
 *   data-parallel exchange (all-reduce it over the ranks, e.g. NCCL, then prb_learner_apply with world_size);
 *   td_mse_host [M] (may be NULL): mean squared TD error per node.
 * prb_learner_apply: Adam on the nodes whose activity count is positive.  prb_learner_sync_targets: update_target /
 *   sync_neighbor_target_q_network (learner.py:190-199,256-270).  prb_learner_read: 0 weights, 1 target weights, 2 exp_avg,
 *   3 exp_avg_sq ([M * P] floats), 4 steps ([M]). */
int prb_replay_sample_all (prb_env *env, int32_t batch, uint64_t seed, int64_t min_rows, float *obs_dev, int32_t *action_dev,
                           float *reward_dev, float *next_obs_dev, float *done_dev, float *weights_dev, float *active_dev);
int prb_learner_init (prb_env *env, float lr, float gamma, float beta1, float beta2, float adam_eps, int target_mode);
int prb_learner_sync_targets (prb_env *env);
int prb_learner_grad (prb_env *env, int32_t batch, const float *obs_dev, const int32_t *action_dev, const float *reward_dev,
                      const float *next_obs_dev, const float *done_dev, const float *weights_dev, const float *active_dev,
                      float *bucket_dev, float *td_mse_host);
int prb_learner_apply (prb_env *env, const float *bucket_dev, int32_t world_size);
int prb_learner_read (prb_env *env, int32_t which, float *out_host, int64_t n_floats);

int prb_counters_host (prb_env *env, prb_counters *out /* [E] */);
int prb_counters_dev (prb_env *env, prb_counters *out_dev /* [E] device */);
int prb_queue_bytes (prb_env *env, int32_t e, int32_t *out /* [L] */);
/* prb_queue_bytes for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same result as prb_queue_bytes; the device-side staging buffer is a file-scope static __device__ array of
 * PRB_NOALLOC_MAX_LINKS int32 (16 KB, held by every GPU that loads the library until exit) instead of a buffer owned
 * by the handle, and the call performs no cudaMalloc / cudaFree / allocating library call.  out: host, at least L
 * entries (out_capacity = its length).  Configurations with more than PRB_NOALLOC_MAX_LINKS switch links are
 * rejected with PRB_EINVAL.  Environments handed over to the serial engine are brought up to date by prb_handovers
 * (which may allocate); while some are pending this call returns PRB_ESTATE.  Every call on a device shares the one
 * array, which keeps the previous call's contents (this call clears the part it uses first): calls that overlap on the
 * GPU (different streams or handles) race; serialise them. */
#define PRB_NOALLOC_MAX_LINKS 4096
#define PRB_NOALLOC_MAX_DEVICES 64
int prb_queue_bytes_noalloc (prb_env *env, int32_t e, int32_t *out /* [>= L] */, int32_t out_capacity);
int prb_sync (prb_env *env);
/* number of kernel launches issued by this handle so far */
int64_t prb_launch_count (prb_env *env);

#ifdef __cplusplus
}
#endif
#endif
