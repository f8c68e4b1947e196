"""The CUDA decode kernel's float parser (csrc/float_parse.cuh, host+device code) fuzzed on the CPU against glibc strtof:
random floats at 9/17/19 digits, exact rounding midpoints and their neighbours, random decimal strings with exponents,
> 19-digit strings (must match or be refused), grammar corner cases, and every token of the reference CSV."""
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parse_fuzz_against_strtof(tmp_path, golden_dir):
    exe = str(tmp_path / "parse_fuzz")
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "helpers", "parse_fuzz.cpp")], check=True)
    toks = np.load(os.path.join(golden_dir, "csv_tokens.npz"))["tokens"]
    tf = tmp_path / "toks.txt"
    tf.write_text("\n".join(toks.tolist()) + "\n")
    r = subprocess.run([exe, "300000", str(tf)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "mismatches=0" in r.stdout


def test_integer_tokens_against_strtof(tmp_path):
    """The at-most-8-digit integer shortcut of the CSV / JSON number parser (tyre pressures, firmware version) vs glibc strtof."""
    exe = str(tmp_path / "int_token_fuzz")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-o", exe,
                    os.path.join(ROOT, "tests", "helpers", "int_token_fuzz.cpp")], check=True)
    r = subprocess.run([exe, "1000000"], capture_output=True, text=True)
    assert r.returncode == 0 and "0 mismatches" in r.stdout, r.stdout[-500:]
