"""Experiment (not applied): event-time slots per node so that select is a minimum over 16-byte loads."""
def edit(p, pairs):
    s=open(p).read()
    for a,b in pairs:
        assert s.count(a)==1,(p,a[:70],s.count(a)); s=s.replace(a,b)
    open(p,'w').write(s)
R='/root/repo/'
edit(R+'prisma_b200/csrc/par_core.cuh', [
("""  int32_t DEG;               /* row length of in_links / out_links */""",
 """  int32_t DEG;               /* row length of in_links / out_links */
  int32_t S;                 /* event-time slots per node: DEG + 1 Receive slots (in-links, access link), DEG + 1
                                TransmitComplete slots (out-links, access link), padded to a multiple of 4 */"""),
("""  const int32_t *flow_begin; /* [N+1] flows with source x: [flow_begin[x], flow_begin[x+1]) */""",
 """  const int32_t *flow_begin; /* [N+1] flows with source x: [flow_begin[x], flow_begin[x+1]) */
  const int16_t *rslot;      /* [LA] slot (index into ev_ts) of the Receive of link l at its receiving node */
  const int16_t *tslot;      /* [LA] slot of the TransmitComplete of link l at its sending node */"""),
("""      off_ping_index, off_pol_pos, off_fmin_f, off_head, off_mid, off_tail, off_s_head, off_s_tail, off_fly_szc, off_fly_ord, off_cand;""",
 """      off_ping_index, off_pol_pos, off_fmin_f, off_head, off_mid, off_tail, off_s_head, off_s_tail, off_fly_szc, off_fly_ord, off_cand,
      off_ev_ts;"""),
("""  uint32_t *t_next, *fmin_ts, *busy, *w_ts;   /* offsets from sc->base; busy: offset + 1, 0 = idle since before base */""",
 """  uint32_t *t_next, *fmin_ts, *busy, *w_ts;   /* offsets from sc->base; busy: offset + 1, 0 = idle since before base */
  uint32_t *ev_ts;                            /* [N*S] time of the pending Receive / TransmitComplete of every link end */"""),
("""  d.fly_ord = (POrd *) (blob + ps.off_fly_ord);
}""","""  d.fly_ord = (POrd *) (blob + ps.off_fly_ord);
  d.ev_ts = (uint32_t *) (blob + ps.off_ev_ts);
}"""),
("""            d.q_bytes[l] = qb + size;
            return;
          }""","""            d.q_bytes[l] = qb + size;
            d.ev_ts[ps.tslot[l]] = d.busy[l] - 1;     /* a packet waits: the completion is an event */
            return;
          }"""),
("""    d.mid[l] = (uint16_t) (m + 1);
    acc.tx_started++;""","""    d.mid[l] = (uint16_t) (m + 1);
    d.ev_ts[ps.tslot[l]] = (uint16_t) (m + 1) != d.tail[l] ? now + tx : P_FAR32;
    if (d.head[l] == m) d.ev_ts[ps.rslot[l]] = now + tx + delay;       /* the wire was empty: this is its head */
    acc.tx_started++;"""),
("""        if (kind == PK_RCV) d.head[l] = (uint16_t) (i + 1);""","""        if (kind == PK_RCV)
          {
            uint16_t h2 = (uint16_t) (i + 1);
            d.head[l] = h2;
            d.ev_ts[ps.rslot[l]] = h2 != d.mid[l] ? d.w_ts[l * ps.W + (h2 & (ps.W - 1))] : P_FAR32;
          }"""),
("""  /* earliest pending event of node x */
  PRB_HD PCand
  select (int x)
  {
    PCand b;
    b.ts = P_FAR32; b.kind = PK_NONE; b.idx = -1;
    const int W = ps.W;""","""  PRB_HD void
  rebuild_slots (int x)
  {
    const int W = ps.W, la = st.L + x;
    uint32_t *e = d.ev_ts + x * ps.S;
    for (int k = 0; k < ps.S; k++) e[k] = P_FAR32;
    for (int j = 0; j <= ps.DEG; j++)
      {
        int l = j < ps.DEG ? ps.in_links[x * ps.DEG + j] : la;
        if (l >= 0 && d.head[l] != d.mid[l]) e[j] = d.w_ts[l * W + (d.head[l] & (W - 1))];
        l = j < ps.DEG ? ps.out_links[x * ps.DEG + j] : la;
        if (l >= 0 && d.mid[l] != d.tail[l]) e[ps.DEG + 1 + j] = d.busy[l] - 1;
      }
  }

  PRB_HD void
  refresh_node (int x)
  {
    rescan_flows (x);
    rebuild_slots (x);
  }

  PRB_HD PCand
  select (int x)
  {
    struct alignas (16) Quad { uint32_t x, y, z, w; };
    const Quad *e = (const Quad *) (d.ev_ts + x * ps.S);
    uint32_t mt = P_FAR32;
    int mk = -1;
    bool tie_s = false;
    PRB_SELECT_UNROLL
    for (int q = 0; q < ps.S / 4; q++)
      {
        Quad v = e[q];
        uint32_t t4[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
        for (int i = 0; i < 4; i++)
          {
            tie_s = tie_s || t4[i] == mt;
            if (t4[i] < mt)
              {
                mt = t4[i]; mk = 4 * q + i; tie_s = false;
              }
          }
      }
    const int m = st.map_overlay[x];
    const uint32_t tp = m >= 0 ? rel32 (d.t_ping[m], base) : P_FAR32;
    const uint32_t tf = d.fmin_f[x] >= 0 ? d.fmin_ts[x] : P_FAR32;
    uint32_t g = mt < tp ? mt : tp;
    g = tf < g ? tf : g;
    PCand b;
    b.ts = g; b.kind = PK_NONE; b.idx = -1;
    if (g == P_FAR32) return b;
    const int same = (mt == g ? 1 : 0) + (tp == g ? 1 : 0) + (tf == g ? 1 : 0);
    if (PRB_UNLIKELY (same > 1 || (mt == g && tie_s))) return select_ties (x);
    if (tp == g)
      {
        b.kind = PK_PING; b.idx = m;
      }
    else if (tf == g)
      {
        b.kind = PK_FLOW; b.idx = d.fmin_f[x];
      }
    else
      {
        const bool rcv = mk <= ps.DEG;
        const int j = rcv ? mk : mk - ps.DEG - 1;
        b.kind = rcv ? PK_RCV : PK_TXC;
        b.idx = j < ps.DEG ? (rcv ? ps.in_links[x * ps.DEG + j] : ps.out_links[x * ps.DEG + j]) : st.L + x;
      }
    return b;
  }

  PRB_HD PCand
  select_ties (int x)
  {
    PCand b;
    b.ts = P_FAR32; b.kind = PK_NONE; b.idx = -1;
    const int W = ps.W;"""),
("""    for (int f = ps.flow_begin[x]; f < ps.flow_begin[x + 1]; f++) schedule_flow (f, true, 0);
    rescan_flows (x);
  }""","""    for (int f = ps.flow_begin[x]; f < ps.flow_begin[x + 1]; f++) schedule_flow (f, true, 0);
    refresh_node (x);
  }"""),
])
edit(R+'prisma_b200/csrc/prisma_b200.cu', [
("""      for (int x = gl; x < N; x += T) sim.rescan_flows (x);""","""      for (int x = gl; x < N; x += T) sim.refresh_node (x);"""),
("""  place (buf, pm.dcode_txc, o[6]);
  CUDA_TRY (cudaMalloc (&env->d_ptables, buf.size ()));""","""  place (buf, pm.dcode_txc, o[6]); place (buf, pm.rslot, o[7]);
  size_t o_tslot;
  place (buf, pm.tslot, o_tslot);
  CUDA_TRY (cudaMalloc (&env->d_ptables, buf.size ()));"""),
("""  s.dcode_txc = (const uint16_t *) (b + o[6]);
  s.tr_prev = nullptr; s.pol_off = nullptr;
  return PRB_OK;""","""  s.dcode_txc = (const uint16_t *) (b + o[6]);
  s.rslot = (const int16_t *) (b + o[7]); s.tslot = (const int16_t *) (b + o_tslot);
  s.tr_prev = nullptr; s.pol_off = nullptr;
  return PRB_OK;"""),
])
edit(R+'tests/host_harness/host_sim.cpp', [
("""          sim.set_window (until_ns);
          sim.rescan_flows (x);""","""          sim.set_window (until_ns);
          sim.refresh_node (x);"""),
])
edit(R+'prisma_b200/csrc/model.h', [
("""  std::vector<int16_t> in_links, out_links;""","""  std::vector<int16_t> in_links, out_links, rslot, tslot;"""),
("""  ps.off_fly_szc = o; o += LA;
  ps.pblob_bytes = align_up (o, 16);""","""  ps.off_fly_szc = o; o += LA;
  ps.S = align_up (2 * (ps.DEG + 1), 4);
  o = align_up (o, 16);
  ps.off_ev_ts = o; o += 4 * N * ps.S;
  if (N * ps.S > 32767)
    {
      pm.why = "too many link ends";
      return false;
    }
  pm.rslot.assign (LA, 0); pm.tslot.assign (LA, 0);
  for (int x = 0; x < N; x++)
    for (int j = 0; j <= ps.DEG; j++)
      {
        int li = j < ps.DEG ? pm.in_links[(size_t) x * ps.DEG + j] : L + x;
        int lo = j < ps.DEG ? pm.out_links[(size_t) x * ps.DEG + j] : L + x;
        if (li >= 0) pm.rslot[li] = (int16_t) (x * ps.S + j);
        if (lo >= 0) pm.tslot[lo] = (int16_t) (x * ps.S + ps.DEG + 1 + j);
      }
  ps.pblob_bytes = align_up (o, 16);"""),
("""      d.fmin_f[x] = -1;
    }""","""      d.fmin_f[x] = -1;
      for (int k = 0; k < ps.S; k++) d.ev_ts[x * ps.S + k] = P_FAR32;
    }"""),
("""  ps.in_links = pm.in_links.data (); ps.out_links = pm.out_links.data ();""","""  ps.in_links = pm.in_links.data (); ps.out_links = pm.out_links.data ();
  ps.rslot = pm.rslot.data (); ps.tslot = pm.tslot.data ();"""),
])
