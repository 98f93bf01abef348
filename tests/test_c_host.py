"""The C-ABI as a C programmer meets it: include/edison_b200.h must be valid C99 and C++, and examples/host_kws.c -- the
firmware's own loop (app.c:167-221) written against the firmware's entry points -- must link against the library and, on a
GPU, print the oracle's bytes."""
import os
import subprocess

import numpy as np
import pytest

import common

ROOT = common.ROOT
LIBDIR = os.path.join(ROOT, "edison_b200")


def build(tmp_path):
    exe = str(tmp_path / "host_kws")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "host_kws.c"), "-L" + LIBDIR, "-ledison_b200",
                           "-Wl,-rpath," + LIBDIR, "-Wl,--allow-shlib-undefined", "-o", exe])
    return exe


def test_header_is_c99_and_cxx(tmp_path):
    for compiler, std, ext in (("gcc", "-std=c99", "c"), ("g++", "-std=c++11", "cpp")):
        src = tmp_path / ("t." + ext)
        src.write_text('#include "edison_b200.h"\nint main(void) { return EDISON_NET_IN == 403 ? 0 : 1; }\n')
        subprocess.check_call([compiler, std, "-Wall", "-Wextra", "-Werror", "-pedantic", "-fsyntax-only",
                               "-I" + os.path.join(ROOT, "include"), str(src)])


def test_example_links_against_the_library(tmp_path):
    assert os.path.exists(build(tmp_path))


def test_example_fails_loudly_without_a_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("this box has a GPU")
    pcm = tmp_path / "clip.s16le"
    np.zeros(31 * 1024, dtype=np.int16).tofile(pcm)
    r = subprocess.run([build(tmp_path), str(pcm)], capture_output=True, text=True)
    assert r.returncode != 0            # firmware semantics: init failure is fatal (Error_Handler), never a silent CPU result
    assert "net_out" not in r.stdout


@pytest.mark.gpu
def test_example_prints_the_oracle_bytes(tmp_path, oracle):
    clip = common.synth_frames(31, seed=77).reshape(-1)
    pcm = tmp_path / "clip.s16le"
    clip.astype("<i2").tofile(pcm)
    r = subprocess.run([build(tmp_path), str(pcm)], capture_output=True, text=True, check=True)
    lines = dict(l.split(" ", 1) for l in r.stdout.strip().splitlines())
    _, want_sm, want_ni = oracle.kws(clip, 1)
    assert [int(v) for v in lines["net_in"].split()] == want_ni[0].tolist()
    assert [int(v) for v in lines["net_out"].split()] == want_sm[0].tolist()
    assert lines["batched"] == "identical"
    assert lines["keyword"].split()[0] in open(os.path.join(LIBDIR, "data", "keywords.txt")).read()
