"""The cgo host shims (go/icicle_bn254, go/icicle_bls12377) cannot be compiled here (no Go toolchain): what can be checked is
that every C identifier they use is declared by the header they include, that the BLS12-377 package is exactly what
tools/gen_go_bls12_377.py makes of the BN254 one, and that the sizes it hard-codes are the curve's."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _read(*parts):
    with open(os.path.join(ROOT, *parts)) as f:
        return f.read()


def _c_identifiers(pkg):
    used = set()
    d = os.path.join(ROOT, "go", pkg)
    for name in os.listdir(d):
        if name.endswith(".go"):
            used |= set(re.findall(r"\bC\.([gG]16_[A-Za-z0-9_]+)", _read("go", pkg, name)))
    return used


def test_bn254_shim_uses_only_declared_symbols():
    header = _read("include", "g16b200.h")
    used = _c_identifiers("icicle_bn254")
    assert "g16_prove_bn254" in used and "g16_pk_upload" in used
    missing = sorted(s for s in used if not re.search(r"\b" + s + r"\b", header))
    assert not missing, missing


def test_bls12_377_shim_uses_only_declared_symbols():
    header = _read("include", "g16b200_bls12_377.h") + _read("include", "g16b200.h")
    used = _c_identifiers("icicle_bls12377")
    assert "g16_bls12_377_prove" in used and "G16_BLS12_377_PROVE_OUT_BYTES" in used and "G16_CURVE_BLS12_377" in used
    assert not [s for s in used if "bn254" in s.lower()]
    missing = sorted(s for s in used if not re.search(r"\b" + s + r"\b", header))
    assert not missing, missing
    # functions must be the renamed ones: only the shared descriptor types keep the plain prefix
    plain = sorted(s for s in used if s.startswith("g16_") and not s.startswith("g16_bls12_377_"))
    assert plain == ["g16_pk_desc", "g16_pk_file_desc", "g16_stats"], plain


def test_bls12_377_shim_is_up_to_date_and_sized_for_the_curve(tmp_path):
    before = {n: _read("go", "icicle_bls12377", n) for n in os.listdir(os.path.join(ROOT, "go", "icicle_bls12377"))}
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "gen_go_bls12_377.py"), str(tmp_path)], stdout=subprocess.DEVNULL)
    after = {}
    for n in os.listdir(tmp_path):
        with open(tmp_path / n) as f:
            after[n] = f.read()
    assert before == after, "go/icicle_bls12377 is stale: run tools/gen_go_bls12_377.py"
    src = after["icicle.go"]
    assert re.findall(r"&out\[(\d+)\]", src) == ["0", "0", "12", "24", "36", "48"]      # 96-byte G1 points, then the G2 point
    assert "span(96)" in after["dump.go"] and "span(192)" in after["dump.go"] and "span(64)" not in after["dump.go"]
    assert "#define G16_BLS12_377_PROVE_OUT_BYTES 576" in _read("include", "g16b200_bls12_377.h")
