"""CPU-side checks of the boundary: the shared object loads, exports every symbol the
header declares, and its pure-host utilities agree with the numpy statement.
No CUDA compute is called here."""
import os
import re

import numpy as np
import pytest

import __graft_entry__ as entry
from loqa_hub_b200 import _lib, api, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", autouse=True)
def built():
    entry.build()


def test_library_loads_and_exports_every_declared_symbol():
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "gpuaudio.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(ga_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.ga_abi_version() == 1


def test_header_is_plain_c_and_links_from_c(tmp_path):
    """include/gpuaudio.h must be consumable by cgo, i.e. by a C compiler: compile it as strict C99
    and link a C program that takes the address of every declared entry point against the .so."""
    import subprocess
    header = open(os.path.join(ROOT, "include", "gpuaudio.h")).read()
    names = sorted(set(re.findall(r"\b(ga_[a-z0-9_]+)\s*\(", re.sub(r"/\*.*?\*/", "", header, flags=re.S))))
    src = tmp_path / "abi.c"
    src.write_text('#include "gpuaudio.h"\n#include <stdio.h>\nint main(void) {\n  void *p[] = {' +
                   ", ".join("(void *)%s" % n for n in names) +
                   '};\n  printf("%d %d\\n", (int)(sizeof p / sizeof p[0]), ga_abi_version());\n  return 0;\n}\n')
    exe = tmp_path / "abi"
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-Wno-pedantic",
                           "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                           "-L", os.path.join(ROOT, "loqa_hub_b200"), "-lgpuaudio",
                           "-Wl,-rpath," + os.path.join(ROOT, "loqa_hub_b200")])
    out = subprocess.check_output([str(exe)]).decode().split()
    assert int(out[0]) == len(names) == len(_lib.SYMBOLS) and int(out[1]) == 1


def test_no_device_means_a_loud_error_not_a_fallback():
    if api.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(_lib.GpuAudioError) as e:
        api.Context(max_streams=4, max_batch_msgs=16, ring_bytes=1 << 16)
    assert e.value.code == _lib.GA_ERR_CUDA and "no CPU fallback" in str(e.value)


def test_product_does_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "loqa_hub_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower(), os.path.join(dirpath, f)


def test_host_synth_matches_numpy_statement():
    for g, r in [(0, 0), (3, 5), (511, 7), (12, 9)]:
        a = api.synth_host(synth.SEED, g, r, 0, 5000, 80000)
        b = synth.relay_view(g, r, 5000, 80000)
        assert np.array_equal(a, b), (g, r)
    a = api.synth_host(synth.SEED, 7, 2, 79000, 1000, 80000)
    assert np.array_equal(a, synth.relay_view(7, 2, 1000, 80000, first=79000))
    assert 9000 < np.abs(synth.relay_view(0, 0, 80000).astype(np.int32)).max() < 20000


def test_host_batch_layout_and_relay_gain_ordering():
    b = api.synth_batch_host(synth.SEED, 4, 16, 8, 15, 320, n_threads=3).reshape(15, 16, 320)
    assert np.array_equal(b, synth.batch_time_major(4, 16, 8, 15, 320))
    # within a collision group the wake-window energies are strictly ordered by relay gain
    e = (b.astype(np.float64) ** 2).sum(axis=(0, 2)).reshape(2, 8)
    assert (np.diff(e, axis=1) < 0).all()
    assert (e[:, :-1] / e[:, 1:] > 1.05).all()


def test_audio_hash_wire_format():
    """"sha256:" + lowercase hex (API.md:170) — pure host code, needs no device."""
    import hashlib
    from loqa_hub_b200.voice_event import format_audio_hash
    d = hashlib.sha256(b"abc").digest()
    assert format_audio_hash(d) == "sha256:ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"
    assert format_audio_hash(bytes(range(224, 256))) == "sha256:" + bytes(range(224, 256)).hex()
    with pytest.raises(ValueError):
        format_audio_hash(b"short")


def test_reference_arm_of_the_bench_never_loads_the_product_library():
    """`bench.py --impl reference` times the reference's CPU path only: it must print the contract's
    JSON line (same metric / unit / config as the GPU arm, a thread-scaling table) without mapping
    libgpuaudio.so — its input comes from oracle/synth_oracle.c."""
    import json
    import subprocess
    import sys
    code = ("import sys, json; sys.argv = ['bench.py', '--impl', 'reference', '--workload', 'cfg2', '--steps', '1', "
            "'--warmup', '0', '--cpu-threads', '2']; import bench; bench.main(); "
            "sys.stderr.write('MAPS_HAVE_PRODUCT=%d\\n' % ('libgpuaudio' in open('/proc/self/maps').read()))")
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "MAPS_HAVE_PRODUCT=0" in r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "audio-s/s" and line["higher_is_better"] is True
    assert line["config"]["workload"].startswith("cfg2: configs[1]")
    assert set(line["cpu_baseline"]["scaling"]) == {"1", "2"} and line["cpu_baseline"]["kind"] == "port"
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["value"] > 0
