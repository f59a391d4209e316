"""The drop-in binary must not fall through to the reference's CPU loops.

polymps_gpu = the reference's unmodified main.o + host-side objects + polymps_b200/host/pmps_shim.o.  Every engine-class
method main.o references must be (a) defined by the shim — a forward to the C ABI or a loud exit — or (b) on the explicit list
of HOST-side methods (construction, input parsing, parameter set-up, bucket geometry, mesh reading / motion, output) that the
drop-in deliberately leaves to the reference.  A per-particle method in neither set would run on stale host arrays.
CPU-only: needs the reference objects (oracle/_ref/obj, built by __graft_entry__.build() where /root/reference exists)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAIN_O = os.path.join(ROOT, "oracle", "_ref", "obj", "main.o")
HOST = os.path.join(ROOT, "polymps_b200", "host")
SHIM_O = os.path.join(HOST, "_build", "pmps_shim.o")
BIN = os.path.join(HOST, "_build", "polymps_gpu")
MAP = os.path.join(HOST, "_build", "polymps_gpu.map")

pytestmark = pytest.mark.skipif(not (os.path.isfile(MAIN_O) and os.path.isfile(SHIM_O) and os.path.isfile(BIN)),
                                reason="reference objects / drop-in binary not built here (no /root/reference)")

# host-side methods the drop-in leaves to the reference (none of them loops over particles inside the time step)
HOST_SIDE = [
    r"^Mps\w+::Mps\w+\(\)$", r"^PolygonMesh::PolygonMesh\(\)$",                      # constructors
    r"^MpsInputOutput::",                                                             # .json/.grid parsing, .vtu/.pvd writing, console
    r"^MpsParticle::setParameters\(", r"^MpsBucket::allocateMemory\(", r"^MpsPressure::allocateMemory\(",
    r"^MpsBoundaryCondition::setBucketBC\(",                                          # once, before the loop (src/main.cpp:118); periodic domains are refused by pmps_create
    r"^PolygonMesh::(initPolygonMesh|initWijnNeigh|readPolygonMeshFile|writePolygonMeshFile|updateForcedPolygonMesh)\(",
]


def _nm(args):
    return subprocess.run(["nm"] + args, check=True, capture_output=True, text=True).stdout.splitlines()


def _engine_symbols_main_needs():
    out = []
    for ln in _nm(["-u", "-C", MAIN_O]):
        s = ln.strip()[2:].strip() if ln.strip().startswith("U ") else ""
        if re.match(r"^(Mps\w+|PolygonMesh)::", s):
            out.append(s)
    return sorted(set(out))


def _shim_defined(demangle=True):
    out = []
    for ln in _nm(["--defined-only"] + (["-C"] if demangle else []) + [SHIM_O]):
        m = re.match(r"^[0-9a-f]+ T (.*)$", ln)
        if m:
            out.append(m.group(1))
    return sorted(set(out))


def test_every_engine_method_main_reaches_is_forwarded_or_host_side():
    need = _engine_symbols_main_needs(); shim = set(_shim_defined())
    assert len(need) > 60
    fall = [s for s in need if s not in shim and not any(re.search(p, s) for p in HOST_SIDE)]
    assert not fall, "these methods would run the reference's CPU code on stale host arrays: %s" % fall
    # the four the round-1 shim missed are now in it
    for name in ("MpsParticleCollision::checkDynamicParticleCollisions", "MpsShifting::calcShifting", "MpsShifting::calcNormalParticles",
                 "MpsShifting::calcWallNormalParticles", "MpsViscosity::calcWallSlipViscosityInteractionVal", "MpsViscosity::calcWallNoSlipViscosityInteractionVal"):
        assert any(s.startswith(name + "(") for s in shim), name


def test_link_has_one_strong_definition_and_takes_the_shims():
    mk = open(os.path.join(HOST, "Makefile")).read()
    assert "-Wl,--allow-multiple-definition" not in mk.split("_build/polymps_gpu:")[1], "the link must not tolerate duplicate definitions"
    # address range of pmps_shim.o's .text in the link map; every shim symbol of the binary must lie inside it
    rng = None
    lines = open(MAP).read().splitlines()
    for k, ln in enumerate(lines):
        m = re.match(r"^ \.text\s+0x([0-9a-f]+)\s+0x([0-9a-f]+)\s+(\S*pmps_shim\.o)$", ln)
        if not m and ln.strip() == ".text" and k + 1 < len(lines):
            m = re.match(r"^\s+0x([0-9a-f]+)\s+0x([0-9a-f]+)\s+(\S*pmps_shim\.o)$", lines[k + 1])
        if m:
            rng = (int(m.group(1), 16), int(m.group(1), 16) + int(m.group(2), 16)); break
    assert rng, "pmps_shim.o .text not found in the link map"
    mangled = set(_shim_defined(demangle=False))
    checked = 0
    for ln in _nm(["--defined-only", BIN]):
        m = re.match(r"^([0-9a-f]+) [TW] (\S+)$", ln)
        if m and m.group(2) in mangled:
            a = int(m.group(1), 16)
            assert rng[0] <= a < rng[1], "%s resolves outside pmps_shim.o" % m.group(2)
            checked += 1
    assert checked == len(mangled), (checked, len(mangled))
