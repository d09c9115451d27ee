/* TEST INFRASTRUCTURE — harness that turns the reference's own ns-3 program into
 * the parity oracle `oracle/_ref/prisma_ref`.
 *
 * What is compiled into that binary (see oracle/build_ref.py):
 *   - vendored ns-3 modules, from the sources where they lie under /root/reference/ns3-gym/src
 *   - every prisma/ns3/ .cc file on the path, UNMODIFIED, including sim.cc
 *     (its `main` is renamed with -Dmain=prisma_sim_main)
 *   - this file: (a) the in-process replacement of the ZMQ/protobuf bridge
 *     (ns3-gym/src/opengym/model/opengym_interface.cc:154-286,391-402,
 *      opengym_env.cc:68-98), (b) trace sinks that write the event log the
 *     CUDA path is compared with, (c) the agent policy that the Python
 *     `Forwarder` would apply (prisma/source/forwarder.py:135-195).
 *
 * Nothing here re-implements simulator behaviour: queues, links, clocks, RNG,
 * routing and the per-packet managers are the reference's code.
 *
 * Harness options (consumed here, everything else is passed to sim.cc):
 *   --h_out=DIR          output directory (events.csv, decisions.csv, summary.txt)
 *   --h_policy=table|random|trace|zero
 *   --h_policy_file=F    table: M rows of M action indices; trace: one action per line
 *   --h_policy_seed=S    seed of the random policy
 *   --h_log_info=0|1     keep the full info string of every decision
 *   --h_events=0|1       write events.csv (default 1)
 *   --h_kat=F            known-answer mode: evaluate ns-3 Time/DataRate primitives and exit
 */
#include "ns3/core-module.h"
#include "ns3/network-module.h"
#include "ns3/internet-module.h"
#include "ns3/opengym-module.h"
#include "ns3/point-to-point-channel.h"
#include "point-to-point-net-device.h"
#include "my-tag.h"
#include "compute-stats-v2.h"

#include <cstdio>
#include <unistd.h>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <fstream>
#include <sstream>

using namespace ns3;

int prisma_sim_main (int argc, char *argv[]);

namespace {

struct HarnessCfg
{
  std::string out = ".";
  std::string policy = "zero";
  std::string policyFile;
  uint64_t policySeed = 1;
  bool logInfo = false;
  bool events = true;
  std::string kat;
} g_cfg;

FILE *g_ev = 0;
FILE *g_dec = 0;
bool g_hooked = false;
uint32_t g_nSwitch = 0;
int g_nIfaces = 0;
std::vector<std::vector<int> > g_table;
std::vector<int> g_actionTrace;
size_t g_actionPos = 0;
uint64_t g_rng = 0;
uint64_t g_nDecisions = 0, g_nSteps = 0;
bool g_summaryDone = false;
unsigned long g_nHops = 0, g_nDataHops = 0;   /* Receive() on switch<->switch devices, tagged packets */

struct LastRx
{
  int64_t t; uint32_t node, dev; uint64_t uid; uint32_t size;
  uint32_t type, src, dst, lastHop, nextHop, valable;
} g_lastRx;

uint64_t
NextRand ()
{
  /* xorshift64*: only used to drive the `random` policy, logged so it can be replayed */
  g_rng ^= g_rng >> 12; g_rng ^= g_rng << 25; g_rng ^= g_rng >> 27;
  return g_rng * 2685821657736338717ULL;
}

void
ParseCtx (const std::string &ctx, uint32_t &node, uint32_t &dev)
{
  /* "/NodeList/<n>/DeviceList/<d>/..." */
  node = dev = 0;
  sscanf (ctx.c_str (), "/NodeList/%u/DeviceList/%u/", &node, &dev);
}

bool
ReadTag (Ptr<const Packet> p, MyTag &t)
{
  return p->PeekPacketTag (t);
}

uint32_t
QueueBytes (uint32_t node, uint32_t dev)
{
  Ptr<PointToPointNetDevice> d = DynamicCast<PointToPointNetDevice> (NodeList::GetNode (node)->GetDevice (dev));
  return d->GetQueue ()->GetNBytes ();
}

void
LogEvent (const char *kind, const std::string &ctx, Ptr<const Packet> p, bool withQueue)
{
  if (!g_ev) return;
  uint32_t node, dev;
  ParseCtx (ctx, node, dev);
  MyTag t;
  bool has = ReadTag (p, t);
  /* kind,t_ns,node,dev,uid,size,type,src,dst,lastHop,nextHop,valable,rejected,qbytes */
  fprintf (g_ev, "%s,%ld,%u,%u,%lu,%u,%d,%d,%d,%d,%d,%d,%d,%d\n", kind,
           (long) Simulator::Now ().GetNanoSeconds (), node, dev, (unsigned long) p->GetUid (), p->GetSize (),
           has ? (int) t.GetSimpleValue () : -1, has ? (int) t.GetSource () : -1,
           has ? (int) t.GetFinalDestination () : -1, has ? (int) t.GetLastHop () : -1,
           has ? (int) t.GetNextHop () : -1, has ? (int) t.GetTrafficValable () : -1,
           has ? (int) t.GetRejectedPacket () : -1,
           withQueue ? (int) QueueBytes (node, dev) : -1);
}

/* traffic node, device 1: a source hands a packet to its access link (PoissonGeneratorApplication::SendPacket) */
void OnMacTx (std::string ctx, Ptr<const Packet> p)
{
  uint32_t node, dev; ParseCtx (ctx, node, dev);
  if (node >= g_nSwitch && dev == 1) LogEvent ("INJ", ctx, p, false);
}
void OnEnqueue (std::string ctx, Ptr<const Packet> p)
{
  uint32_t node, dev; ParseCtx (ctx, node, dev);
  if (node < g_nSwitch && dev >= 2) LogEvent ("ENQ", ctx, p, true);
}
void OnQueueDrop (std::string ctx, Ptr<const Packet> p)
{
  uint32_t node, dev; ParseCtx (ctx, node, dev);
  if (node < g_nSwitch && dev >= 2) LogEvent ("DRP", ctx, p, true);
}
void OnPhyTxBegin (std::string ctx, Ptr<const Packet> p)
{
  uint32_t node, dev; ParseCtx (ctx, node, dev);
  if (node < g_nSwitch && dev >= 2) LogEvent ("TXB", ctx, p, true);
}
void OnMacRx (std::string ctx, Ptr<const Packet> p)
{
  uint32_t node, dev; ParseCtx (ctx, node, dev);
  if (node >= g_nSwitch) return;
  MyTag t;
  bool has = ReadTag (p, t);
  g_lastRx.t = Simulator::Now ().GetNanoSeconds ();
  g_lastRx.node = node; g_lastRx.dev = dev; g_lastRx.uid = p->GetUid (); g_lastRx.size = p->GetSize ();
  g_lastRx.type = has ? t.GetSimpleValue () : 255;
  g_lastRx.src = has ? t.GetSource () : 0; g_lastRx.dst = has ? t.GetFinalDestination () : 0;
  g_lastRx.lastHop = has ? t.GetLastHop () : 0; g_lastRx.nextHop = has ? t.GetNextHop () : 0;
  g_lastRx.valable = has ? t.GetTrafficValable () : 0;
  if (dev >= 2 && has && t.GetSimpleValue () <= 4)
    {
      g_nHops++;
      if (t.GetSimpleValue () == 0) g_nDataHops++;
    }
  LogEvent (dev >= 2 ? "RCV" : "ARR", ctx, p, false);
}

void
InstallHooks ()
{
  if (g_hooked) return;
  g_hooked = true;
  g_nSwitch = NodeList::GetNNodes () / 2;   /* sim.cc:327-331: N switch nodes then N traffic nodes */
  std::string evName = g_cfg.out + "/events.csv";
  std::string decName = g_cfg.out + "/decisions.csv";
  if (g_cfg.events)
    {
      g_ev = fopen (evName.c_str (), "w");
      fprintf (g_ev, "kind,t_ns,node,dev,uid,size,type,src,dst,lastHop,nextHop,valable,rejected,qbytes\n");
    }
  g_dec = fopen (decName.c_str (), "w");
  fprintf (g_dec, "t_ns,iface,node,uid,ptype,done,reward,action,nobs,obs\n");
  const std::string base = "/NodeList/*/DeviceList/*/$ns3::PointToPointNetDevice/";
  Config::Connect (base + "MacTx", MakeCallback (&OnMacTx));
  Config::Connect (base + "MacRx", MakeCallback (&OnMacRx));
  Config::Connect (base + "PhyTxBegin", MakeCallback (&OnPhyTxBegin));
  Config::Connect (base + "TxQueue/Enqueue", MakeCallback (&OnEnqueue));
  Config::Connect (base + "TxQueue/Drop", MakeCallback (&OnQueueDrop));

  if (g_cfg.policy == "table")
    {
      std::ifstream f (g_cfg.policyFile.c_str ());
      std::string line;
      while (std::getline (f, line))
        {
          std::istringstream iss (line);
          std::vector<int> row; int v;
          while (iss >> v) row.push_back (v);
          if (!row.empty ()) g_table.push_back (row);
        }
      if (g_table.empty ()) { fprintf (stderr, "ref_harness: empty policy table %s\n", g_cfg.policyFile.c_str ()); exit (2); }
    }
  else if (g_cfg.policy == "trace")
    {
      std::ifstream f (g_cfg.policyFile.c_str ());
      int v;
      while (f >> v) g_actionTrace.push_back (v);
    }
  g_rng = g_cfg.policySeed * 0x9E3779B97F4A7C15ULL + 0x1234567ULL;
  if (g_rng == 0) g_rng = 1;
}

void
WriteSummary ()
{
  if (g_summaryDone) return;
  g_summaryDone = true;
  std::string name = g_cfg.out + "/summary.txt";
  FILE *f = fopen (name.c_str (), "w");
  ComputeStats cs;
  fprintf (f, "n_switch %u\n", g_nSwitch);
  fprintf (f, "n_ifaces %d\n", g_nIfaces);
  fprintf (f, "end_ns %ld\n", (long) Simulator::Now ().GetNanoSeconds ());
  fprintf (f, "decisions %lu\n", (unsigned long) g_nDecisions);
  fprintf (f, "env_steps %lu\n", (unsigned long) g_nSteps);
  fprintf (f, "packet_hops %lu\n", g_nHops);
  fprintf (f, "data_hops %lu\n", g_nDataHops);
  fprintf (f, "overlay_injected %d\n", cs.getGlobalOverlayPacketsInjected ());
  fprintf (f, "overlay_arrived %d\n", cs.getGlobalOverlayPacketsArrived ());
  fprintf (f, "overlay_lost %d\n", cs.getGlobalOverlayPacketsLost ());
  fprintf (f, "underlay_injected %d\n", cs.getGlobalUnderlayPacketsInjected ());
  fprintf (f, "underlay_arrived %d\n", cs.getGlobalUnderlayPacketsArrived ());
  fprintf (f, "underlay_lost %d\n", cs.getGlobalUnderlayPacketsLost ());
  float v;
  uint32_t bits;
  v = cs.getAverage (cs.getGlobalE2eDelay ()); memcpy (&bits, &v, 4);
  fprintf (f, "avg_e2e_delay %.9g 0x%08x\n", v, bits);
  v = cs.getAverage (cs.getGlobalCost ()); memcpy (&bits, &v, 4);
  fprintf (f, "avg_cost %.9g 0x%08x\n", v, bits);
  v = cs.getAverage (cs.getGlobalUnderlayE2eDelay ()); memcpy (&bits, &v, 4);
  fprintf (f, "avg_underlay_e2e_delay %.9g 0x%08x\n", v, bits);
  v = cs.getAverage (cs.getGlobalUnderlayCost ()); memcpy (&bits, &v, 4);
  fprintf (f, "avg_underlay_cost %.9g 0x%08x\n", v, bits);
  v = cs.getSignalingOverhead (); memcpy (&bits, &v, 4);
  fprintf (f, "signaling_overhead %.9g 0x%08x\n", v, bits);
  fprintf (f, "n_e2e %lu\n", (unsigned long) cs.getGlobalE2eDelay ().size ());
  fprintf (f, "n_cost %lu\n", (unsigned long) cs.getGlobalCost ().size ());

  /* device -> peer map of every switch node (device 0 = loopback, 1 = access link) */
  for (uint32_t n = 0; n < g_nSwitch; n++)
    {
      Ptr<Node> node = NodeList::GetNode (n);
      for (uint32_t d = 1; d < node->GetNDevices (); d++)
        {
          Ptr<PointToPointNetDevice> dev = DynamicCast<PointToPointNetDevice> (node->GetDevice (d));
          Ptr<PointToPointChannel> ch = DynamicCast<PointToPointChannel> (dev->GetChannel ());
          Ptr<NetDevice> other = (ch->GetDevice (0) == dev) ? ch->GetDevice (1) : ch->GetDevice (0);
          fprintf (f, "dev %u %u peer %u\n", n, d, other->GetNode ()->GetId ());
        }
    }
  /* global-routing next hops exactly as PRISMA looks them up (data-packet-manager.cc:186-190,279-283) */
  for (uint32_t n = 0; n < g_nSwitch; n++)
    {
      Ptr<Node> node = NodeList::GetNode (n);
      Ptr<Ipv4> ipv4 = node->GetObject<Ipv4> ();
      Ptr<Ipv4RoutingProtocol> routing = ipv4->GetRoutingProtocol ();
      for (uint32_t d = 0; d < g_nSwitch; d++)
        {
          std::string ip = "10.2.2." + std::to_string (d + 1);
          Ipv4Header h;
          h.SetDestination (Ipv4Address (ip.c_str ()));
          Ptr<Packet> pkt = Create<Packet> (20);
          Socket::SocketErrno err;
          Ptr<Ipv4Route> r = routing->RouteOutput (pkt, h, 0, err);
          int ifi = -1;
          if (r && r->GetOutputDevice ()) ifi = r->GetOutputDevice ()->GetIfIndex ();
          fprintf (f, "route %u %u dev %d\n", n, d, ifi);
        }
    }
  fclose (f);
  if (g_ev) { fclose (g_ev); g_ev = 0; }
  if (g_dec) { fclose (g_dec); g_dec = 0; }
}

int
RunKat (const std::string &file)
{
  /* Each input line is one query; the answer is computed by ns-3's own code.
   *   tx <rate-string> <bytes>     -> DataRate::CalculateBytesTxTime in ns      (data-rate.cc:235-240)
   *   s2ns <hex double bits>       -> Seconds(double).GetNanoSeconds()          (nstime.h:456-475)
   *   ns2s <int64 ns>              -> NanoSeconds(ns).GetSeconds() as hex bits  (nstime.h:505-523)
   *   rate <rate-string>           -> DataRate(str).GetBitRate()
   *   exp <seed> <run> <stream> <mean-hex> -> first ExponentialRandomVariable draw (hex bits)
   */
  std::ifstream f (file.c_str ());
  std::string line;
  while (std::getline (f, line))
    {
      std::istringstream iss (line);
      std::string op; iss >> op;
      if (op == "tx")
        {
          std::string rate; uint32_t bytes; iss >> rate >> bytes;
          printf ("tx %s %u %ld\n", rate.c_str (), bytes, (long) DataRate (rate).CalculateBytesTxTime (bytes).GetNanoSeconds ());
        }
      else if (op == "txbps")
        {
          uint64_t bps; uint32_t bytes; iss >> bps >> bytes;
          printf ("txbps %lu %u %ld\n", (unsigned long) bps, bytes, (long) DataRate (bps).CalculateBytesTxTime (bytes).GetNanoSeconds ());
        }
      else if (op == "s2ns")
        {
          std::string hx; iss >> hx;
          uint64_t bits = strtoull (hx.c_str (), 0, 16); double d; memcpy (&d, &bits, 8);
          printf ("s2ns %s %ld\n", hx.c_str (), (long) Seconds (d).GetNanoSeconds ());
        }
      else if (op == "ns2s")
        {
          long ns; iss >> ns;
          double d = NanoSeconds (ns).GetSeconds (); uint64_t bits; memcpy (&bits, &d, 8);
          printf ("ns2s %ld %016lx\n", ns, (unsigned long) bits);
        }
      else if (op == "rate")
        {
          std::string rate; iss >> rate;
          printf ("rate %s %lu\n", rate.c_str (), (unsigned long) DataRate (rate).GetBitRate ());
        }
    }
  return 0;
}

} // namespace

namespace ns3 {

NS_OBJECT_ENSURE_REGISTERED (OpenGymSpace);
NS_OBJECT_ENSURE_REGISTERED (OpenGymDiscreteSpace);
NS_OBJECT_ENSURE_REGISTERED (OpenGymBoxSpace);
NS_OBJECT_ENSURE_REGISTERED (OpenGymDataContainer);
NS_OBJECT_ENSURE_REGISTERED (OpenGymDiscreteContainer);
NS_OBJECT_ENSURE_REGISTERED (OpenGymEnv);
NS_OBJECT_ENSURE_REGISTERED (OpenGymInterface);

TypeId OpenGymSpace::GetTypeId ()
{
  static TypeId tid = TypeId ("ns3::OpenGymSpace").SetParent<Object> ().SetGroupName ("OpenGym");
  return tid;
}
TypeId OpenGymDiscreteSpace::GetTypeId ()
{
  static TypeId tid = TypeId ("ns3::OpenGymDiscreteSpace").SetParent<OpenGymSpace> ().SetGroupName ("OpenGym")
    .AddConstructor<OpenGymDiscreteSpace> ();
  return tid;
}
TypeId OpenGymBoxSpace::GetTypeId ()
{
  static TypeId tid = TypeId ("ns3::OpenGymBoxSpace").SetParent<OpenGymSpace> ().SetGroupName ("OpenGym")
    .AddConstructor<OpenGymBoxSpace> ();
  return tid;
}
TypeId OpenGymDataContainer::GetTypeId ()
{
  static TypeId tid = TypeId ("ns3::OpenGymDataContainer").SetParent<Object> ().SetGroupName ("OpenGym");
  return tid;
}
TypeId OpenGymDiscreteContainer::GetTypeId ()
{
  static TypeId tid = TypeId ("ns3::OpenGymDiscreteContainer").SetParent<OpenGymDataContainer> ().SetGroupName ("OpenGym")
    .AddConstructor<OpenGymDiscreteContainer> ();
  return tid;
}
TypeId OpenGymEnv::GetTypeId ()
{
  static TypeId tid = TypeId ("ns3::OpenGymEnv").SetParent<Object> ().SetGroupName ("OpenGym");
  return tid;
}
TypeId OpenGymInterface::GetTypeId ()
{
  static TypeId tid = TypeId ("ns3::OpenGymInterface").SetParent<Object> ().SetGroupName ("OpenGym")
    .AddConstructor<OpenGymInterface> ();
  return tid;
}

void OpenGymEnv::SetOpenGymInterface (Ptr<OpenGymInterface> i) { m_openGymInterface = i; }
void OpenGymEnv::Notify () { if (m_openGymInterface) m_openGymInterface->Notify (this); }
void OpenGymEnv::NotifySimulationEnd () { if (m_openGymInterface) m_openGymInterface->NotifySimulationEnd (); }

OpenGymInterface::OpenGymInterface (uint32_t port) : m_port (port), m_init (false), m_index (0), m_k (0)
{
  InstallHooks ();
  m_index = g_nIfaces++;
}

/* One Gym step, in the order of OpenGymInterface::NotifyCurrentState
 * (opengym_interface.cc:213-286): Init() once (queries both spaces), then
 * observation, reward, game-over, info, [transport], ExecuteActions. */
void
OpenGymInterface::Notify (Ptr<OpenGymEnv> env)
{
  if (!m_init)
    {
      m_init = true;
      env->GetObservationSpace ();
      Ptr<OpenGymDiscreteSpace> as = DynamicCast<OpenGymDiscreteSpace> (env->GetActionSpace ());
      m_k = as ? as->GetN () : 0;
    }
  Ptr<OpenGymDataContainer> obsC = env->GetObservation ();
  float reward = env->GetReward ();
  bool done = env->GetGameOver ();
  std::string info = env->GetExtraInfo ();
  std::vector<int64_t> obs = obsC->AsInt64 ();

  /* Forwarder.step (forwarder.py:149-152): action 0 at the destination, on the
   * reset/train-step sentinel (-1) and on control packets (1000); else the policy. */
  int action = 0;
  bool sentinel = obs.empty () || obs[0] == -1 || obs[0] == 1000;
  bool policyStep = !sentinel && obs[0] != m_index && m_k > 0;
  if (policyStep)
    {
      if (g_cfg.policy == "table")
        action = g_table.at (m_index).at ((size_t) obs[0]);
      else if (g_cfg.policy == "random")
        action = (int) (NextRand () % (uint64_t) m_k);
      else if (g_cfg.policy == "trace")
        {
          if (g_actionPos >= g_actionTrace.size ()) { fprintf (stderr, "ref_harness: action trace exhausted\n"); exit (3); }
          action = g_actionTrace[g_actionPos++];
        }
      g_nDecisions++;
    }
  bool initStep = obs.size () == 1 && obs[0] == -1;
  if (!initStep) g_nSteps++;
  if (g_dec)
    {
      uint32_t rbits; memcpy (&rbits, &reward, 4);
      fprintf (g_dec, "%ld,%d,%u,%lu,%d,%d,%08x,%d,%lu", (long) Simulator::Now ().GetNanoSeconds (), m_index,
               initStep ? 0u : g_lastRx.node, initStep ? 0ul : (unsigned long) g_lastRx.uid,
               initStep ? -1 : (int) g_lastRx.type, done ? 1 : 0, rbits, policyStep ? action : -1,
               (unsigned long) obs.size ());
      for (size_t i = 0; i < obs.size (); i++) fprintf (g_dec, ",%ld", (long) obs[i]);
      if (g_cfg.logInfo) fprintf (g_dec, ",\"%s\"", info.c_str ());
      fprintf (g_dec, "\n");
    }
  Ptr<OpenGymDiscreteContainer> act = CreateObject<OpenGymDiscreteContainer> ((uint32_t) m_k);
  act->SetValue ((uint32_t) action);
  env->ExecuteActions (act);
}

void
OpenGymInterface::NotifySimulationEnd ()
{
  WriteSummary ();
}

} // namespace ns3

int
main (int argc, char *argv[])
{
  std::vector<char *> pass;
  pass.push_back (argv[0]);
  for (int i = 1; i < argc; i++)
    {
      std::string a (argv[i]);
      if (a.compare (0, 8, "--h_out=") == 0) g_cfg.out = a.substr (8);
      else if (a.compare (0, 11, "--h_policy=") == 0) g_cfg.policy = a.substr (11);
      else if (a.compare (0, 16, "--h_policy_file=") == 0) g_cfg.policyFile = a.substr (16);
      else if (a.compare (0, 16, "--h_policy_seed=") == 0) g_cfg.policySeed = strtoull (a.substr (16).c_str (), 0, 10);
      else if (a.compare (0, 13, "--h_log_info=") == 0) g_cfg.logInfo = atoi (a.substr (13).c_str ()) != 0;
      else if (a.compare (0, 11, "--h_events=") == 0) g_cfg.events = atoi (a.substr (11).c_str ()) != 0;
      else if (a.compare (0, 8, "--h_kat=") == 0) g_cfg.kat = a.substr (8);
      else pass.push_back (argv[i]);
    }
  if (!g_cfg.kat.empty ())
    {
      int rc = RunKat (g_cfg.kat);
      fflush (stdout);
      _exit (rc);   /* skip ns-3's static destructors: nothing was simulated */
    }
  int rc = prisma_sim_main ((int) pass.size (), pass.data ());
  return rc;
}
