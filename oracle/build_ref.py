#!/usr/bin/env python3
"""TEST INFRASTRUCTURE — builds oracle/_ref/prisma_ref, the reference's own ns-3 program.

This is NOT the reference's build system (waf is not run: it needs Python <= 3.10,
protoc, libzmq and libprotobuf, none of which exist here).  It is a direct g++
recipe over the sources where they lie under /root/reference:

  * vendored ns-3 modules core, stats, network, bridge, traffic-control, internet,
    point-to-point (channel + ppp header only), applications (packet sink only)
  * prisma/ns3_model/ipv4-interface.cc in place of ns-3's own copy (what
    prisma/source/run_ns3.py:30 does with rsync)
  * every prisma/ns3/*.cc that sim.cc links (conf-loader, ospf, tunnel-nd,
    tcp-application and subnet are dead code and left out)
  * oracle/ref_harness/ (our in-process opengym stand-in + trace sinks)

Outputs go to oracle/_ref/ only (git-ignored; travels to the GPU box).  No
reference source is copied: headers are exposed through symlinks in
oracle/_ref/include/ns3, exactly like waf's build/ns3 directory.
"""
import os
import re
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("PRB_REFERENCE", "/root/reference")
NS3 = os.path.join(REF, "ns3-gym", "src")
PRISMA = os.path.join(REF, "prisma", "ns3")
OUT = os.path.join(HERE, "_ref")
INC = os.path.join(OUT, "include")
OBJ = os.path.join(OUT, "obj")
HARNESS = os.path.join(HERE, "ref_harness")

# modules whose headers are exposed (sim.cc includes their *-module.h)
HEADER_MODULES = ["core", "stats", "network", "bridge", "traffic-control", "internet",
                  "point-to-point", "applications", "csma", "mobility", "mpi",
                  "virtual-net-device", "config-store"]
# modules whose sources are compiled in full (model/ helper/ utils/)
FULL_MODULES = ["core", "stats", "network", "bridge", "traffic-control", "internet", "mpi"]

EXCLUDE = {
    # platform / optional-feature sources that waf only adds conditionally (src/core/wscript:268-312)
    "core": {"win32-system-wall-clock-ms.cc", "int64x64-cairo.cc", "int64x64-double.cc",
             "cairo-wideint.c", "realtime-simulator-impl.cc", "wall-clock-synchronizer.cc",
             "example-as-test.cc"},
    "stats": {"sqlite-data-output.cc"},
    # the reference links its own DropTailQueue / Ipv4Interface instead of these
    "network": {"drop-tail-queue.cc"},
    "internet": {"ipv4-interface.cc"},
    "bridge": set(), "traffic-control": set(), "mpi": set(),
}
EXTRA_NS3 = [
    "point-to-point/model/point-to-point-channel.cc",
    "point-to-point/model/point-to-point-remote-channel.cc",
    "point-to-point/model/ppp-header.cc",
    "applications/model/packet-sink.cc",
    "applications/helper/packet-sink-helper.cc",
]
PRISMA_SRC = [
    "sim.cc", "point-to-point-net-device.cc", "point-to-point-helper.cc", "drop-tail-queue.cc",
    "poisson-application.cc", "poisson-app-helper.cc", "big-signaling-application.cc",
    "big-signaling-app-helper.cc", "my-tag.cc", "compute-stats-v2.cc", "packet-routing-gym.cc",
    "packet-manager.cc", "data-packet-manager.cc", "ping-forward-packet-manager.cc",
    "ping-back-packet-manager.cc", "small-signaling-packet-manager.cc",
    "big-signaling-packet-manager.cc",
    # referenced only by the unused ScheduleHelloMessages (sim.cc:1050-1053), needed to link
    "ipv4-ospf-routing.cc", "ospf-tag.cc", "conf-loader.cc", "subnet.cc",
]

CXX = os.environ.get("CXX", "g++")
CXXFLAGS = ["-std=gnu++11", "-O2", "-g0", "-fPIC", "-w", "-pthread",
            "-DNS3_BUILD_PROFILE_OPTIMIZED", "-DHAVE_SYS_STAT_H", "-DHAVE_DIRENT_H",
            "-DHAVE_STDLIB_H", "-DHAVE_GETENV", "-DHAVE_SIGNAL_H", "-DHAVE_PTHREAD_H",
            "-DINT64X64_USE_128", "-DHAVE___UINT128_T", "-DHAVE_STDINT_H", "-DHAVE_INTTYPES_H",
            "-DHAVE_SYS_TYPES_H"]


def sh(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return r.returncode, r.stdout


def symlink(src, dst):
    if os.path.islink(dst) or os.path.exists(dst):
        os.remove(dst)
    os.symlink(src, dst)


def module_dirs(mod):
    return [d for d in (os.path.join(NS3, mod, s) for s in ("model", "helper", "utils"))
            if os.path.isdir(d)]


def public_headers(mod):
    """Header list of a module = the 'headers.source' entries of its wscript (read, not run)."""
    txt = open(os.path.join(NS3, mod, "wscript")).read()
    m = re.search(r"(?<![a-z])headers\.source\s*=\s*\[(.*?)\]", txt, re.S)
    names = re.findall(r"'([^']+\.h)'", m.group(1)) if m else []
    return [n for n in names if os.path.exists(os.path.join(NS3, mod, n))]


def make_includes():
    ns3inc = os.path.join(INC, "ns3")
    os.makedirs(ns3inc, exist_ok=True)
    for mod in HEADER_MODULES:
        for d in module_dirs(mod):
            for h in os.listdir(d):
                if h.endswith(".h"):
                    symlink(os.path.join(d, h), os.path.join(ns3inc, h))
        hdrs = public_headers(mod)
        if mod == "core":
            hdrs.append("model/int64x64-128.h")
        with open(os.path.join(ns3inc, mod + "-module.h"), "w") as f:
            guard = "NS3_MODULE_" + mod.upper().replace("-", "_")
            f.write("#ifndef %s\n#define %s\n" % (guard, guard))
            for h in hdrs:
                f.write('#include "ns3/%s"\n' % os.path.basename(h))
            f.write("#endif\n")
    # generated by waf's configure step in the reference build (src/core/wscript:157)
    with open(os.path.join(ns3inc, "core-config.h"), "w") as f:
        f.write("#ifndef NS3_CORE_CONFIG_H\n#define NS3_CORE_CONFIG_H\n#endif\n")
    # sim.cc:66 includes netanim but instantiates nothing from it
    with open(os.path.join(ns3inc, "netanim-module.h"), "w") as f:
        f.write("/* unused by sim.cc; intentionally empty */\n")
    # the reference's forks win over the vendored headers of the same name
    for h in ("point-to-point-net-device.h", "point-to-point-helper.h", "drop-tail-queue.h"):
        symlink(os.path.join(PRISMA, h), os.path.join(ns3inc, h))
    symlink(os.path.join(HARNESS, "ns3", "opengym-module.h"), os.path.join(ns3inc, "opengym-module.h"))


def shadow(src, extra_headers):
    """Compile `src` from a directory of symlinks so that its quoted includes of
    forked headers resolve to the reference's forks (same class layout everywhere)."""
    d = os.path.join(OUT, "shadow", os.path.basename(src).replace(".", "_"))
    os.makedirs(d, exist_ok=True)
    srcdir = os.path.dirname(src)
    for h in os.listdir(srcdir):
        if h.endswith(".h"):
            symlink(os.path.join(srcdir, h), os.path.join(d, h))
    for h in extra_headers:
        symlink(h, os.path.join(d, os.path.basename(h)))
    dst = os.path.join(d, os.path.basename(src))
    symlink(src, dst)
    return dst


def jobs():
    out = []
    for mod in FULL_MODULES:
        for d in module_dirs(mod):
            for s in sorted(os.listdir(d)):
                if s.endswith(".cc") and s not in EXCLUDE.get(mod, set()):
                    out.append((os.path.join(d, s), mod + "__" + s, []))
    fork_hdrs = [os.path.join(PRISMA, h) for h in ("point-to-point-net-device.h", "compute-stats-v2.h")]
    for s in EXTRA_NS3:
        p = os.path.join(NS3, s)
        if s.startswith("point-to-point/"):
            p = shadow(p, fork_hdrs)
        out.append((p, s.replace("/", "__"), []))
    out.append((os.path.join(REF, "prisma", "ns3_model", "ipv4-interface.cc"), "prisma_model__ipv4-interface.cc",
                ["-I" + os.path.join(NS3, "internet", "model")]))
    for s in PRISMA_SRC:
        extra = ["-include", "cassert"]
        if s == "sim.cc":
            extra += ["-Dmain=prisma_sim_main"]
        out.append((os.path.join(PRISMA, s), "prisma__" + s, extra))
    out.append((os.path.join(HARNESS, "ref_harness.cc"), "harness__ref_harness.cc", []))
    out.append((os.path.join(HARNESS, "log_components.cc"), "harness__log_components.cc", []))
    return out


def compile_one(job):
    src, name, extra = job
    obj = os.path.join(OBJ, name[:-3].replace(".", "_") + ".o")
    if os.path.exists(obj) and os.path.getmtime(obj) > os.path.getmtime(os.path.realpath(src)):
        return obj, 0, ""
    cmd = [CXX] + CXXFLAGS + ["-I" + INC, "-I" + PRISMA, "-I" + os.path.join(HARNESS)] + extra + ["-c", src, "-o", obj]
    rc, log = sh(cmd)
    return obj, rc, log


def main():
    if not os.path.isdir(NS3):
        print("build_ref: %s not present; keeping prebuilt oracle/_ref if any" % NS3)
        return 0
    os.makedirs(OBJ, exist_ok=True)
    make_includes()
    js = jobs()
    objs, failed = [], []
    with ThreadPoolExecutor(max_workers=int(os.environ.get("PRB_JOBS", os.cpu_count() or 4))) as ex:
        for obj, rc, log in ex.map(compile_one, js):
            objs.append(obj)
            if rc != 0:
                failed.append((obj, log))
    for obj, log in failed:
        print("FAILED", obj)
        print(log[-3000:])
    if failed:
        return 1
    exe = os.path.join(OUT, "prisma_ref")
    rc, log = sh([CXX, "-O2", "-pthread", "-o", exe] + objs + ["-lrt", "-ldl"])
    if rc != 0:
        print(log[-6000:])
        return 1
    print("built", exe)
    return 0


if __name__ == "__main__":
    sys.exit(main())
