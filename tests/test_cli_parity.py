"""the command line of the renderer (csrc/host/main.cpp -> rtw_cli) against the
reference's RestOfLife/main.cpp:41-143 — same flags, same clamps, same warning texts, same P3 output.

CPU: rtw_cli prints its argument diagnostics before it asks for a device, so the texts can be
checked here without a GPU — against the strings of main.cpp:66-137 written out below AND, when
oracle/_ref/ref_cli exists (the reference's own main.cpp compiled for the host), against what the
reference's program really prints for the same arguments.
GPU: the clamped sizes end up in the PPM header, and the PPM of `-s 0 -ns 1` is byte for byte the
oracle's frame through Director::printPPM's formula (Director.cpp:1010-1031)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "raytracing-weekend_b200", "rtw_cli")
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "ref_cli")
REF_ROOT = "/root/reference/RestOfLife"

# (arguments, the diagnostics main.cpp prints for them: main.cpp:66-137)
CASES = [
    (["-s", "7"], ["WARNING: Scene number 7 out of range. Maximum scene number: 4", "WARNING: Using a scene value of 4"]),
    (["-s", "-1"], ["WARNING: Scene number -1 out of range. Maximum scene number: 4", "WARNING: Using a scene value of 4"]),
    (["-s", "abc"], ["Invalid scene number: abc"]),
    (["-ns", "0"], ["WARNING: Number of samples 0 is out of range. Using a value of 20"]),
    (["-ns", "20000"], ["WARNING: Number of samples 20000 is out of range. Using a value of 10240"]),
    (["-ns", "x"], ["Invalid number of samples: x"]),
    (["-dx", "100"], ["WARNING: Width (-dx) 100 out of range. Using a value of 320"]),
    (["-dx", "5000"], ["WARNING: Width (-dx) 5000 out of range. Using a value of 3840"]),
    (["-dx", "q"], ["Invalid image width (-dx): q"]),
    (["-dy", "10"], ["WARNING: Width (-dy) 10 out of range. Using a value of 200"]),          # (sic) main.cpp:131
    (["-dy", "9000"], ["WARNING: Width (-dy) 9000 out of range. Using a value of 2240"]),
    (["-dy", "z"], ["Invalid image height (-dy): z"]),
    (["-s", "4", "-ns", "5", "-dx", "320", "-dy", "200"], []),
    (["-dx", "100", "-dy", "10", "-ns", "0", "-s", "9"],
     ["WARNING: Scene number 9 out of range. Maximum scene number: 4", "WARNING: Using a scene value of 4",
      "WARNING: Number of samples 0 is out of range. Using a value of 20",
      "WARNING: Width (-dx) 100 out of range. Using a value of 320", "WARNING: Width (-dy) 10 out of range. Using a value of 200"]),
]


def _diagnostics(stderr_text):
    return [ln.rstrip() for ln in stderr_text.splitlines() if ln.startswith("WARNING:") or ln.startswith("Invalid ")]


def _need_cli():
    if not os.path.exists(CLI):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "raytracing-weekend_b200"), "rtw_cli"])


@pytest.mark.parametrize("argv,expected", CASES)
def test_argument_diagnostics_are_the_references(argv, expected):
    _need_cli()
    # without a device rtw_cli stops at rtw_create (exit 1), AFTER the argument diagnostics; -d 99 makes sure of it anywhere
    out = subprocess.run([CLI] + argv + ["-d", "99"], capture_output=True, text=True)
    assert _diagnostics(out.stderr) == expected
    assert "rtw_create failed" in out.stderr and out.returncode != 0     # no CPU fallback


@pytest.mark.parametrize("argv,expected", [c for c in CASES if c[0][0] != "-s" or c[0][1] not in "0123"])
def test_reference_program_prints_the_same_diagnostics(argv, expected):
    if not (os.path.exists(REF_CLI) and os.path.isdir(REF_ROOT)):
        pytest.skip("oracle/_ref/ref_cli or /root/reference not present")
    env = dict(os.environ, REF_SKIP_LAUNCH="1")
    out = subprocess.run([REF_CLI] + argv, cwd=REF_ROOT, env=env, capture_output=True, text=True)
    assert out.returncode == 0
    assert _diagnostics(out.stderr) == expected
    # and the sizes it settled on (the PPM header, Director.cpp:1013)
    toks = out.stdout.split(None, 4)
    want_x = {"100": "320", "5000": "3840", "320": "320"}.get(dict(zip(argv[::2], argv[1::2])).get("-dx", ""), "1200")
    want_y = {"10": "200", "9000": "2240", "200": "200"}.get(dict(zip(argv[::2], argv[1::2])).get("-dy", ""), "600")
    assert toks[:4] == ["P3", want_x, want_y, "255"]


def test_help_text_lists_the_reference_flags():
    _need_cli()
    out = subprocess.run([CLI, "-h"], capture_output=True, text=True)
    assert out.returncode == 0
    for line in ("-s N           Scene Selection number N (N: 0, 1, 2, etc.)", "-ns N          Sample each pixel N times (N: 1, 2, etc.)",
                 "-dx Nx         Output image width (x dimension)", "-dy Ny         Output image height (y dimension)",
                 "-h             This help message.", "-v             Verbose output.", "-g             Debug output."):
        assert line in out.stderr
    if os.path.exists(REF_CLI):
        ref = subprocess.run([REF_CLI, "-h"], capture_output=True, text=True)
        ref_lines = [ln for ln in ref.stderr.splitlines() if ln.strip().startswith("-")]
        ours = out.stderr.splitlines()
        assert ref_lines and all(ln in ours for ln in ref_lines)


def _ppm_of(frame):
    """Director::printPPM (Director.cpp:1010-1031): top row first, int(255.99f * clamp(c, 0, 1))."""
    c = np.clip(frame[::-1, :, :3], np.float32(0), np.float32(1))
    v = (np.float32(255.99) * c).astype(np.int32)
    head = "P3\n%d %d\n255\n" % (frame.shape[1], frame.shape[0])
    return head + "".join("%d %d %d\n" % tuple(px) for px in v.reshape(-1, 3))


@pytest.mark.gpu
def test_cli_ppm_is_the_oracle_frame(rtw, orc, tmp_path):
    out = tmp_path / "c.ppm"
    r = subprocess.run([CLI, "-s", "0", "-ns", "1", "-dx", "320", "-dy", "200", "-o", str(out), "-v"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "INFO: Output image dimensions: 320x200" in r.stderr and "INFO: Number of rays sent per pixel: 1" in r.stderr
    assert "INFO: Scene number selected: 0" in r.stderr and "INFO: Scene description: Cornell box" in r.stderr
    s = rtw.Scene(0, 320, 200)
    sums, _ = orc.OracleScene(s.desc, 13).render(320, 200, 1, mode=0)
    assert out.read_text() == _ppm_of(orc.tonemap(sums, 1))


@pytest.mark.gpu
def test_cli_clamps_reach_the_image(tmp_path):
    out = tmp_path / "d.ppm"
    r = subprocess.run([CLI, "-s", "3", "-ns", "1", "-dx", "100", "-dy", "10", "-o", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert _diagnostics(r.stderr) == ["WARNING: Width (-dx) 100 out of range. Using a value of 320",
                                      "WARNING: Width (-dy) 10 out of range. Using a value of 200"]
    assert out.read_text().split()[:4] == ["P3", "320", "200", "255"]
    assert "INFO: Took " in r.stderr
