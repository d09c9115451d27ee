/* prisma_b200 — host-side model builder: turns a prb_config into the read-only tables, the state-blob
 * layout and the initial state of one environment (what sim.cc:274-652 builds as ns-3 objects).
 * Plain C++ (no CUDA): used by prisma_b200.cu, and by the test-only host harness.
 */
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "sim_core.cuh"
#include "par_core.cuh"

namespace prb {

/* ---- ns-3 Time / DataRate arithmetic on the host ---------------------------------------------------- */

/* Seconds (double) -> Time in ns: the long double constructor of int64x64_t splits the value, rounds the
 * fraction to 64 bits (+0.5), Time::From multiplies by 1e9 and keeps the high word
 * (ns3-gym/src/core/model/int64x64-128.h:94-126, nstime.h:235-236,456-475).  x86-64 only (80-bit x87). */
inline int64_t
host_seconds_to_ns (double s)
{
  typedef unsigned __int128 u128;
  bool neg = s < 0;
  long double v = neg ? -(long double) s : (long double) s;
  long double ip;
  long double fr = modfl (v, &ip);
  const long double two64 = 18446744073709551616.0L;
  fr = fr * two64 + 0.5L;
  u128 hi = (u128) ip;
  uint64_t lo = (uint64_t) fr;
  if (fr >= two64) hi += 1;
  u128 mid = (u128) lo * 1000000000u;
  u128 q_hi = hi * 1000000000u + (mid >> 64);
  uint64_t q_lo = (uint64_t) mid;
  if (!neg) return (int64_t) q_hi;
  return q_lo != 0 ? -(int64_t) q_hi - 1 : -(int64_t) q_hi;
}

/* int64x64_t::Invert (1e9) (int64x64-128.cc:216-232): the Q64.64 reciprocal with 64 more fraction bits,
 * i.e. floor (2^128 / 1e9) by long division, bumped by one when the round trip does not give 1. */
inline void
host_invert_1e9 (uint64_t *hi, uint64_t *lo)
{
  typedef unsigned __int128 u128;
  const u128 v = 1000000000u;
  u128 one = (u128) 1 << 64;
  u128 q1 = one / v, r1 = one % v;
  u128 q0 = (r1 << 64) / v;
  u128 r = (q1 << 64) + q0;
  u128 rh = r >> 64, rl = r & 0xFFFFFFFFFFFFFFFFull;
  u128 tmp = v * rh + ((v * rl) >> 64);
  if ((uint64_t) (tmp >> 64) != 1) r += 1;
  *hi = (uint64_t) (r >> 64);
  *lo = (uint64_t) r;
}

/* DataRate::CalculateBytesTxTime (ns3-gym/src/network/utils/data-rate.cc:235-240) */
inline int64_t
host_tx_time_ns (uint32_t bytes, uint64_t bps)
{
  double s = (double) bytes * 8 / (double) bps;
  return host_seconds_to_ns (s);
}

inline int32_t
host_build_links (int32_t n, const uint8_t *adj, int32_t *from, int32_t *to)
{
  int32_t l = 0;
  for (int i = 0; i < n; i++)
    for (int j = i + 1; j < n; j++)
      if (adj[i * n + j])
        {
          if (from) { from[l] = i; to[l] = j; from[l + 1] = j; to[l + 1] = i; }
          l += 2;
        }
  return l;
}

/* Ipv4GlobalRouting as PRISMA uses it (PopulateRoutingTables, sim.cc:683; LookupGlobal without random
 * ECMP returns the first host route, ipv4-global-routing.cc:139-235; equal-cost exits are installed in
 * next-hop address order = device creation order, global-route-manager-impl.cc:321-333,1342-1440):
 * among the unit-cost shortest paths the first hop with the smallest node id is taken. */
inline int32_t
host_global_routes (int32_t n, const uint8_t *adj, int32_t *route_out)
{
  int32_t nl = host_build_links (n, adj, nullptr, nullptr);
  std::vector<int32_t> from (nl + 1), to (nl + 1), link_of ((size_t) n * n, -1);
  host_build_links (n, adj, from.data (), to.data ());
  for (int l = 0; l < nl; l++) link_of[(size_t) from[l] * n + to[l]] = l;
  std::vector<int32_t> dist (n), fifo (n);
  for (int dst = 0; dst < n; dst++)
    {
      std::fill (dist.begin (), dist.end (), -1);
      int head = 0, tail = 0;
      dist[dst] = 0;
      fifo[tail++] = dst;
      while (head < tail)
        {
          int v = fifo[head++];
          for (int w = 0; w < n; w++)
            if (w != v && link_of[(size_t) v * n + w] >= 0 && dist[w] < 0)
              {
                dist[w] = dist[v] + 1;
                fifo[tail++] = w;
              }
        }
      for (int x = 0; x < n; x++)
        {
          int32_t best = -1;
          if (x != dst && dist[x] > 0)
            for (int y = 0; y < n && best < 0; y++)
              if (y != x && link_of[(size_t) x * n + y] >= 0 && dist[y] == dist[x] - 1)
                best = link_of[(size_t) x * n + y];
          route_out[(size_t) x * n + dst] = best;
        }
    }
  return nl;
}

/* ---- model ------------------------------------------------------------------------------------------ */

struct HostModel
{
  Static st;                       /* table pointers are HOST pointers into the vectors below */
  std::vector<int16_t> link_from, link_to, route, map_overlay, o2u, agent_nbr, flow_src, flow_dst;
  std::vector<uint8_t> link_rc, agent_k, agent_sign_szc;
  std::vector<int64_t> tx_ns;
  std::vector<uint16_t> szc_bytes;
  std::vector<double> flow_mean_s, opt_rejected;
  std::vector<int32_t> flow_of, flow_first;      /* flow_of: [N*N] flow index or -1 (host only) */
  std::vector<int32_t> big_flow_of;              /* [N*N] NN-segment source index or -1 (host only) */
  std::vector<char> blob0;                       /* initial state of one environment */
  std::string error;
};

inline int32_t
align_up (int32_t v, int32_t a)
{
  return (v + a - 1) / a * a;
}

inline bool
build_model (const prb_config &c, HostModel &hm)
{
  Static &st = hm.st;
  memset (&st, 0, sizeof (st));
  const int N = c.n_nodes, M = c.n_overlay;
  if (N < 2 || N > PRB_MAX_NODES || M < 1 || M > N || c.num_envs < 1)
    {
      hm.error = "n_nodes must be 2..250, 1 <= n_overlay <= n_nodes, num_envs >= 1";
      return false;
    }
  if (!c.phys_adj || !c.overlay_adj || !c.map_overlay || !c.traffic_bps)
    {
      hm.error = "phys_adj, overlay_adj, map_overlay and traffic_bps are required";
      return false;
    }
  if (c.link_cap_bps == 0 || c.ping_interval_ns <= 0 || c.sim_time_ns <= 0)
    {
      hm.error = "link_cap_bps, ping_interval_ns and sim_time_ns must be positive";
      return false;
    }
  const int L = host_build_links (N, c.phys_adj, nullptr, nullptr);
  const int LA = L + N;
  if (LA >= 32000)
    {
      hm.error = "too many links";
      return false;
    }
  std::vector<int32_t> from (L + 1), to (L + 1);
  host_build_links (N, c.phys_adj, from.data (), to.data ());
  hm.link_from.assign (LA, 0); hm.link_to.assign (LA, 0); hm.link_rc.assign (LA, 0);
  for (int l = 0; l < L; l++)
    {
      hm.link_from[l] = (int16_t) from[l];
      hm.link_to[l] = (int16_t) to[l];
    }
  /* degrees count the 1s of a full row (sim.cc:346-356); access link rate 1e6*LinkRate*deg (sim.cc:403) */
  std::vector<int> degree (N, 0);
  for (int i = 0; i < N; i++)
    for (int j = 0; j < N; j++)
      if (c.phys_adj[i * N + j] == 1) degree[i]++;
  std::vector<uint64_t> rates;
  rates.push_back (c.link_cap_bps);
  for (int i = 0; i < N; i++)
    {
      uint64_t r = 1000000ull * c.link_cap_bps * (uint64_t) degree[i];
      size_t k = std::find (rates.begin () + 1, rates.end (), r) - rates.begin ();
      if (k == rates.size ()) rates.push_back (r);
      if (k > 255)
        {
          hm.error = "too many distinct access-link rates";
          return false;
        }
      hm.link_from[L + i] = (int16_t) (N + i);
      hm.link_to[L + i] = (int16_t) i;
      hm.link_rc[L + i] = (uint8_t) k;
    }
  /* routes */
  std::vector<int32_t> route ((size_t) N * N);
  if (c.route_out) memcpy (route.data (), c.route_out, sizeof (int32_t) * N * N);
  else host_global_routes (N, c.phys_adj, route.data ());
  hm.route.resize ((size_t) N * N);
  for (size_t i = 0; i < route.size (); i++) hm.route[i] = (int16_t) route[i];
  /* overlay */
  hm.map_overlay.resize (N);
  hm.o2u.assign (M, 0);
  for (int i = 0; i < N; i++)
    {
      hm.map_overlay[i] = (int16_t) c.map_overlay[i];
      if (c.map_overlay[i] >= M)
        {
          hm.error = "map_overlay entry out of range";
          return false;
        }
      if (c.map_overlay[i] >= 0) hm.o2u[c.map_overlay[i]] = (int16_t) i;   /* sim.cc:462-467 */
    }
  std::vector<std::vector<int>> nbr (M);
  for (int i = 0; i < M; i++)                                               /* sim.cc:469-476 */
    for (int j = i; j < M; j++)
      if (c.overlay_adj[i * M + j] == 1)
        {
          nbr[i].push_back (hm.o2u[j]);
          nbr[j].push_back (hm.o2u[i]);
        }
  int K = 1;
  for (int m = 0; m < M; m++) K = std::max (K, (int) nbr[m].size ());
  if (K > PRB_MAX_K)
    {
      hm.error = "overlay degree above PRB_MAX_K";
      return false;
    }
  hm.agent_k.assign (M, 0);
  hm.agent_nbr.assign ((size_t) M * K, 0);
  for (int m = 0; m < M; m++)
    {
      hm.agent_k[m] = (uint8_t) nbr[m].size ();
      for (size_t t = 0; t < nbr[m].size (); t++)
        {
          hm.agent_nbr[(size_t) m * K + t] = (int16_t) nbr[m][t];
          if (route[(size_t) hm.o2u[m] * N + nbr[m][t]] < 0)
            {
              hm.error = "an overlay neighbour is unreachable in the underlay";
              return false;
            }
        }
    }
  /* wire sizes: payload + UDP 8 + IPv4 20 + PPP 2 (data-packet-manager.cc:269-275,365-400) */
  hm.szc_bytes.assign (MAX_SZC, 0);
  hm.szc_bytes[SZC_DATA_INJ] = (uint16_t) (std::max<uint32_t> (c.packet_size, 200) + 30);   /* poisson-application.cc:277-278 */
  hm.szc_bytes[SZC_DATA_FWD] = (uint16_t) (c.packet_size + 30);
  hm.szc_bytes[SZC_PING] = 38;
  int n_szc = SZC_SIGN0;
  hm.agent_sign_szc.assign (M, SZC_SIGN0);
  for (int m = 0; m < M; m++)
    {
      int deg = 0;
      for (int j = 0; j < M; j++)
        if (c.overlay_adj[m * M + j] == 1) deg++;
      uint32_t sz = 0;                                                       /* sim.cc:372-393 */
      if (c.signaling_type == 1) sz = 8 + 8 * (deg + 1);
      else if (c.signaling_type == 2) sz = 24;
      else if (c.signaling_type == 3) sz = 8 + 8 * (2 * deg + 1);
      uint16_t wire = (uint16_t) (sz + 30);
      int k = SZC_SIGN0;
      while (k < n_szc && hm.szc_bytes[k] != wire) k++;
      if (k == n_szc)
        {
          if (n_szc == MAX_SZC)
            {
              hm.error = "too many distinct signalling sizes";
              return false;
            }
          hm.szc_bytes[n_szc++] = wire;
        }
      hm.agent_sign_szc[m] = (uint8_t) k;
    }
  hm.tx_ns.assign (rates.size () * MAX_SZC, 0);
  for (size_t r = 0; r < rates.size (); r++)
    for (int k = 0; k < n_szc; k++)
      hm.tx_ns[r * MAX_SZC + k] = host_tx_time_ns (hm.szc_bytes[k], rates[r]);
  /* flows in creation order (sim.cc:599-632) */
  hm.flow_of.assign ((size_t) N * N, -1);
  uint32_t pkt = std::max<uint32_t> (c.packet_size, 200);
  for (int i = 0; i < N; i++)
    for (int j = 0; j < N; j++)
      {
        bool both = c.map_overlay[i] >= 0 && c.map_overlay[j] >= 0;
        bool active = both || c.activate_underlay_traffic;
        if (i != j && active && c.traffic_bps[i * N + j] > 0)
          {
            hm.flow_of[(size_t) i * N + j] = (int32_t) hm.flow_src.size ();
            hm.flow_src.push_back ((int16_t) i);
            hm.flow_dst.push_back ((int16_t) j);
            hm.flow_mean_s.push_back ((double) (pkt * 8) / (double) c.traffic_bps[i * N + j]);
            if (route[(size_t) i * N + j] < 0)
              {
                hm.error = "a traffic-matrix entry joins disconnected nodes";
                return false;
              }
          }
      }
  const int FP = (int) hm.flow_src.size ();
  /* NN-weight segment sources (sim.cc:634-647): one per directed overlay edge that carries traffic, created next
   * to the Poisson source of the same pair.  They are kept after the Poisson sources; among themselves they keep
   * the (i, j) creation order, which is the order of their simultaneous SendPacket events. */
  hm.big_flow_of.assign ((size_t) N * N, -1);
  int big_szc = 0;
  if (c.train && c.signaling && c.signaling_type == 1 && c.big_signaling_size >= 512 && c.sync_step > 0)
    {
      for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++)
          {
            int mi = c.map_overlay[i], mj = c.map_overlay[j];
            if (hm.flow_of[(size_t) i * N + j] < 0 || mi < 0 || mj < 0 || c.overlay_adj[mi * M + mj] != 1) continue;
            hm.big_flow_of[(size_t) i * N + j] = (int32_t) hm.flow_src.size ();
            hm.flow_src.push_back ((int16_t) i);
            hm.flow_dst.push_back ((int16_t) j);
            hm.flow_mean_s.push_back (0.0);
          }
      /* big-signaling-application.cc:247-250: uint32 * 8 / float -> float, widened; uint32 512 * 8 / double */
      float sync = (float) c.sync_step;
      double rate = (double) ((float) (uint32_t) (c.big_signaling_size * 8u) / sync);
      st.big_gap_ns = host_seconds_to_ns ((double) (512u * 8u) / rate);
      st.big_start_ns = host_seconds_to_ns (0.0001);                        /* AppStartTime, sim.cc:244,644 */
      st.big_segs = (int32_t) (c.big_signaling_size / 512u);                /* :243 */
      const uint16_t wire = 512 + 30;
      big_szc = SZC_SIGN0;
      while (big_szc < n_szc && hm.szc_bytes[big_szc] != wire) big_szc++;
      if (big_szc == n_szc)
        {
          if (n_szc == MAX_SZC)
            {
              hm.error = "too many distinct signalling sizes";
              return false;
            }
          hm.szc_bytes[n_szc++] = wire;
          for (size_t r = 0; r < rates.size (); r++)
            hm.tx_ns[r * MAX_SZC + big_szc] = host_tx_time_ns (wire, rates[r]);
        }
      if (st.big_gap_ns <= 0)
        {
          hm.error = "sync_step / bigSignalingSize give a zero gap between NN segments";
          return false;
        }
    }
  st.big_szc = big_szc;
  if (st.big_segs < 1) st.big_segs = 1;
  const int F = (int) hm.flow_src.size ();
  st.FP = FP;
  hm.flow_first.assign (std::max (F, 1), -1);
  if (c.opt_rejected) hm.opt_rejected.assign (c.opt_rejected, c.opt_rejected + (size_t) N * N);

  st.q_hidden = 2; st.q_pad_ = 0;
  st.E = c.num_envs; st.N = N; st.M = M; st.L = L; st.LA = LA; st.F = F; st.K = K;
  st.S = 2 * LA + M + F;
  st.NK = 2 * LA + M;
  st.NKP = align_up (st.NK, 32);
  st.MA = (int32_t) std::max<uint32_t> (c.moving_average_size, 1);
  st.ma_size = c.moving_average_size;
  st.n_szc = n_szc; st.n_rc = (int32_t) rates.size ();
  st.link_delay_ns = c.link_delay_ns; st.sim_time_ns = c.sim_time_ns; st.ping_interval_ns = c.ping_interval_ns;
  st.max_out_buffer = c.max_out_buffer_size;
  st.ping_as_obs = c.ping_as_obs; st.train = c.train; st.opt_agent = c.opt_agent;
  st.inject_mode = PRB_INJECT_PHILOX; st.policy = PRB_POLICY_ZERO; st.mode = MODE_RUN;
  st.loss_penalty = (float) c.loss_penalty;
  st.seed = c.seed; st.env_id_base = c.env_id_base; st.policy_seed = 1;
  host_invert_1e9 (&st.inv1e9_hi, &st.inv1e9_lo);
  if (st.inv1e9_hi != INV1E9_HI || st.inv1e9_lo != INV1E9_LO)
    {
      hm.error = "int64x64 inverse of 1e9 differs from the constant compiled into the kernels";
      return false;
    }
  /* packet pool: what the byte-limited queues can hold as data packets, plus the wires and slack */
  int P = c.pool_capacity;
  if (P <= 0)
    {
      int64_t per_link = c.max_out_buffer_size / std::max<int> (hm.szc_bytes[SZC_DATA_FWD], 38) + 8;
      if (c.train) per_link *= 2;
      int64_t p = (int64_t) L * per_link + 8ll * N + 64;
      P = (int) std::min<int64_t> (p, 65000);
    }
  if (P > 65000 || P < 8)
    {
      hm.error = "pool_capacity must be 8..65000";
      return false;
    }
  st.P = P;
  /* blob layout: 8-byte arrays, then 4, 2, 1 */
  int32_t o = 0;
  st.off_scal = o; o += align_up ((int32_t) sizeof (Scal), 8);
  st.off_t_ts = o; o += 8 * st.S;
  st.off_sent_mask = o; o += 8 * M * K;
  st.off_t_key = o; o += 4 * st.NKP;
  st.off_t_seq = o; o += 4 * st.S;
  st.FG = (std::max (F, 1) + 31) / 32;
  o = align_up (o, 8);
  st.off_fg_ts = o; o += 8 * st.FG;
  st.off_fg_seq = o; o += 4 * st.FG;
  st.off_fg_slot = o; o += 4 * st.FG;
  st.off_q_bytes = o; o += 4 * LA;
  st.off_f_pos = o; o += 4 * std::max (F, 1);
  st.off_ping_index = o; o += 4 * M;
  st.off_win = o; o += 4 * M * K * st.MA;
  st.off_lost = o; o += 4 * M * PRB_MAX_LOST;
  st.off_n_lost = o; o += 4 * M;
  st.off_q_pkts = o; o += 2 * LA;
  st.off_q_head = o; o += 2 * LA;
  st.off_q_tail = o; o += 2 * LA;
  st.off_w_head = o; o += 2 * LA;
  st.off_w_tail = o; o += 2 * LA;
  st.off_last_dst = o; o += 2 * M;
  st.off_n_win = o; o += M * K;
  st.off_sent_old = o; o += M * K;
  st.blob_bytes = align_up (o, 16);
  /* initial state */
  hm.blob0.assign (st.blob_bytes, 0);
  Dyn d;
  bind_state (st, hm.blob0.data (), d);
  for (int s = 0; s < st.S; s++) d.t_ts[s] = T_IDLE;
  for (int g = 0; g < st.FG; g++)
    {
      d.fg_ts[g] = T_IDLE; d.fg_seq[g] = 0xFFFFFFFFu; d.fg_slot[g] = -1;
    }
  for (int s = 0; s < st.NKP; s++) d.t_key[s] = KEY_IDLE;
  for (int l = 0; l < LA; l++) d.q_head[l] = d.q_tail[l] = d.w_head[l] = d.w_tail[l] = NIL;
  for (int m = 0; m < M; m++) d.last_dst[m] = -1;
  d.sc->free_head = 0;
  /* host pointers */
  st.link_from = hm.link_from.data (); st.link_to = hm.link_to.data (); st.link_rc = hm.link_rc.data ();
  st.tx_ns = hm.tx_ns.data (); st.szc_bytes = hm.szc_bytes.data (); st.route = hm.route.data ();
  st.map_overlay = hm.map_overlay.data (); st.o2u = hm.o2u.data (); st.agent_k = hm.agent_k.data ();
  st.agent_nbr = hm.agent_nbr.data (); st.agent_sign_szc = hm.agent_sign_szc.data ();
  st.flow_src = hm.flow_src.data (); st.flow_dst = hm.flow_dst.data (); st.flow_mean_s = hm.flow_mean_s.data ();
  st.opt_rejected = hm.opt_rejected.empty () ? nullptr : hm.opt_rejected.data ();
  st.flow_first = hm.flow_first.data ();
  return true;
}

/* per-flow chains of a recorded injection trace (each SendPacket schedules the flow's next one,
 * poisson-application.cc:357) */
struct HostTrace
{
  std::vector<int64_t> t;
  std::vector<uint32_t> uid;
  std::vector<uint8_t> flags;
  std::vector<int32_t> next, prev;
};

inline bool
build_trace (HostModel &hm, int64_t n, const int64_t *t_ns, const int32_t *src, const int32_t *dst, const int64_t *uid,
             const uint8_t *valable, const uint8_t *rejected, HostTrace &tr, const uint8_t *ptype = nullptr)
{
  const int N = hm.st.N;
  tr.t.assign (t_ns, t_ns + n);
  tr.uid.resize (n); tr.flags.resize (n); tr.next.assign (n, -1); tr.prev.assign (n, -1);
  std::fill (hm.flow_first.begin (), hm.flow_first.end (), -1);
  std::vector<int32_t> last (hm.flow_src.size (), -1);
  for (int64_t i = 0; i < n; i++)
    {
      if (src[i] < 0 || src[i] >= N || dst[i] < 0 || dst[i] >= N) return false;
      const bool big = ptype && ptype[i] == PRB_BIG_SIGN;
      if (ptype && ptype[i] != PRB_DATA && !big) return false;
      int f = big ? hm.big_flow_of[(size_t) src[i] * N + dst[i]] : hm.flow_of[(size_t) src[i] * N + dst[i]];
      if (f < 0) return false;
      /* an NN segment carries no uid: its slot holds the ordinal of the segment within its source */
      tr.uid[i] = big ? (last[f] >= 0 ? tr.uid[last[f]] + 1u : 1u) : (uint32_t) uid[i];
      tr.flags[i] = (uint8_t) ((valable[i] ? 1 : 0) | (rejected && rejected[i] ? 2 : 0));
      if (last[f] >= 0) tr.next[last[f]] = (int32_t) i; else hm.flow_first[f] = (int32_t) i;
      tr.prev[i] = last[f];
      last[f] = (int32_t) i;
    }
  return true;
}

/* ---- node-parallel engine (par_core.cuh): tables, ring sizes, state layout ------------------------------ */

struct PModel
{
  PStatic ps;                      /* table pointers are HOST pointers into the vectors below */
  std::vector<int16_t> in_links, out_links;
  std::vector<int32_t> ring_off, flow_begin, pol_off;
  std::vector<uint16_t> ring_mask;
  std::vector<char> pblob0;
  bool supported = false;
  std::string why;                 /* why the configuration has to run on the serial engine */
};

inline int32_t
pow2_ceil (int64_t v)
{
  int32_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

inline bool
build_pmodel (const prb_config &c, const HostModel &hm, PModel &pm)
{
  const Static &st = hm.st;
  PStatic &ps = pm.ps;
  memset (&ps, 0, sizeof (ps));
  const int N = st.N, M = st.M, L = st.L, LA = st.LA, F = st.F;
  pm.supported = false;
  if (st.F > st.FP)
    {
      /* training with NN-weight segments is Gym-mode work: the event-serial engine runs it */
      pm.why = "big-signalling sources run on the event-serial engine";
      return false;
    }
  /* lookahead: the smallest packet that can be put on a switch link */
  int64_t tx_min = INT64_MAX;
  for (int k = 0; k < st.n_szc; k++) tx_min = std::min (tx_min, hm.tx_ns[k]);
  if (tx_min <= 0 || st.link_delay_ns < 0)
    {
      pm.why = "zero transmission time";
      return false;
    }
  ps.la_ns = st.link_delay_ns + tx_min;
  /* times inside a launch are 32-bit offsets from its start: the launch may cover 2^32 ns minus the longest
   * transmission + delay that can be scheduled from inside it */
  int64_t tx_max = 0;
  for (size_t i = 0; i < hm.tx_ns.size (); i++) tx_max = std::max (tx_max, hm.tx_ns[i]);
  int64_t reach = tx_max + st.link_delay_ns;
  if (reach > 1000000000ll || ps.la_ns > 1000000000ll)
    {
      pm.why = "a transmission takes longer than a second";
      return false;
    }
  ps.la32 = (uint32_t) ps.la_ns;
  ps.max_chunk_ns = 4294967295ll - 2 * reach - 2 * ps.la_ns - 16;
  /* packets between TransmitStart and the peer's Receive, as seen by a sender that runs up to one lookahead
   * ahead of the receiver: 3 + 2*delay/tx_min */
  int64_t w = 3 + (2 * st.link_delay_ns + tx_min - 1) / tx_min;
  ps.W = pow2_ceil (w);
  if (ps.W > 16)
    {
      pm.why = "link_delay is many minimum transmission times: the wire rings would not fit";
      return false;
    }
  if (N > 4096)
    {
      pm.why = "too many nodes";
      return false;
    }
  int deg = 1;
  std::vector<std::vector<int>> in (N), out (N);
  for (int l = 0; l < L; l++)
    {
      out[hm.link_from[l]].push_back (l);
      in[hm.link_to[l]].push_back (l);
    }
  for (int x = 0; x < N; x++) deg = std::max<int> (deg, (int) std::max (in[x].size (), out[x].size ()));
  ps.DEG = deg;
  pm.in_links.assign ((size_t) N * deg, -1);
  pm.out_links.assign ((size_t) N * deg, -1);
  for (int x = 0; x < N; x++)
    {
      for (size_t j = 0; j < in[x].size (); j++) pm.in_links[(size_t) x * deg + j] = (int16_t) in[x][j];
      for (size_t j = 0; j < out[x].size (); j++) pm.out_links[(size_t) x * deg + j] = (int16_t) out[x][j];
    }
  /* rings: what the byte-limited queue holds as data packets, the wire, and slack for small packets */
  int64_t per = c.max_out_buffer_size / std::max<int> (hm.szc_bytes[SZC_DATA_FWD], 38) + 2 + ps.W;
  if (c.train) per *= 2;
  int32_t rc = std::min (std::max (pow2_ceil (per), 8), 4096);
  /* Lineage retention (par_core.cuh, PRec): consumed records stay readable until the ring wraps, so the rings are
   * longer than the live packets need — PRB_LINEAGE_RECORDS per switch link (default 256), an eighth of it per access
   * link — as long as the rings of all environments stay within PRB_RING_BUDGET_MB (default 24576) of device memory. */
  int32_t lin = 256;
  if (const char *e = getenv ("PRB_LINEAGE_RECORDS")) lin = std::max (8, std::min (pow2_ceil (atoi (e)), 16384));
  int64_t budget = 24576;
  if (const char *e = getenv ("PRB_RING_BUDGET_MB")) budget = std::max<int64_t> (atoll (e), 1);
  auto ring_bytes = [&] (int32_t sw, int32_t ac) { return (int64_t) sizeof (PRec) * ((int64_t) L * sw + (int64_t) (LA - L) * ac) * std::max<int64_t> (c.num_envs, 1); };
  while (lin > rc && ring_bytes (std::max (rc, lin), std::max (8, lin / 8)) > budget * (1ll << 20)) lin /= 2;
  const int32_t rc_sw = std::max (rc, lin), rc_ac = std::max (8, lin / 8);
  ps.walk_max = P_WALK_MAX;
  if (const char *e = getenv ("PRB_LINEAGE_WALK_MAX")) ps.walk_max = std::max (0, atoi (e));    /* tests: 0 = no walk, every such tie is handed over */
  pm.ring_off.assign (LA, 0);
  pm.ring_mask.assign (LA, 0);
  int32_t off = 0;
  for (int l = 0; l < LA; l++)
    {
      int32_t cap = l < L ? rc_sw : rc_ac;
      pm.ring_off[l] = off;
      pm.ring_mask[l] = (uint16_t) (cap - 1);
      off += cap;
    }
  ps.ring_records = off;
  pm.flow_begin.assign (N + 1, 0);
  for (int f = 0; f < F; f++) pm.flow_begin[hm.flow_src[f] + 1]++;
  for (int x = 0; x < N; x++) pm.flow_begin[x + 1] += pm.flow_begin[x];
  for (int f = 1; f < F; f++)
    if (hm.flow_src[f] < hm.flow_src[f - 1])
      {
        pm.why = "flows are not grouped by source";
        return false;
      }
  pm.pol_off.assign (M + 1, 0);
  /* state layout */
  int32_t o = 0;
  ps.off_scal = o; o += align_up ((int32_t) sizeof (PScal), 8);
  ps.off_t_ping = o; o += 8 * M;
  ps.off_f_ts = o; o += 8 * std::max (F, 1);
  ps.off_t_next = o; o += 4 * N;
  ps.off_fmin_ts = o; o += 4 * N;
  ps.off_busy = o; o += 4 * LA;
  ps.off_w_ts = o; o += 4 * LA * ps.W;
  ps.off_sring = o; o += (int32_t) sizeof (PStat) * N * P_SR;
  ps.off_q_bytes = o; o += 4 * LA;
  ps.off_f_pos = o; o += 4 * std::max (F, 1);
  ps.off_ping_index = o; o += 4 * M;
  ps.off_pol_pos = o; o += 4 * M;
  ps.off_fmin_f = o; o += 4 * N;
  ps.off_cand = o; o += 4 * N;
  ps.off_head = o; o += 2 * LA;
  ps.off_mid = o; o += 2 * LA;
  ps.off_tail = o; o += 2 * LA;
  ps.off_s_head = o; o += N;
  ps.off_s_tail = o; o += N;
  ps.off_fly_szc = o; o += LA;
  ps.pblob_bytes = align_up (o, 16);
  pm.pblob0.assign (ps.pblob_bytes, 0);
  PDyn d;
  p_bind_state (ps, pm.pblob0.data (), d);
  for (int x = 0; x < N; x++)
    {
      d.t_next[x] = P_FAR32;
      d.fmin_ts[x] = P_FAR32;
      d.fmin_f[x] = -1;
    }
  for (int m = 0; m < M; m++) d.t_ping[m] = P_FAR;
  for (int f = 0; f < std::max (F, 1); f++) d.f_ts[f] = P_FAR;
  ps.in_links = pm.in_links.data (); ps.out_links = pm.out_links.data ();
  ps.ring_off = pm.ring_off.data (); ps.ring_mask = pm.ring_mask.data ();
  ps.flow_begin = pm.flow_begin.data (); ps.tr_prev = nullptr; ps.pol_off = pm.pol_off.data ();
  pm.supported = true;
  return true;
}

}  // namespace prb
