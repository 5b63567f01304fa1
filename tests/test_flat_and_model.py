"""Contracts of the model and its flat storage."""
import copy
import re
from pathlib import Path

import pytest
import torch

from distributed_neural_network_b200 import Network
from distributed_neural_network_b200.ops import flat as L

ROOT = Path(__file__).resolve().parent.parent


def test_layout_constants_match_cuda_header():
    hdr = (ROOT / "csrc" / "layout.h").read_text()
    vals = {m.group(1): int(m.group(2)) for m in re.finditer(r"constexpr int (k\w+) = (\d+);", hdr)}
    names = ["kOffW1", "kOffB1", "kOffW2", "kOffB2", "kOffF1W", "kOffF1B", "kOffF2W", "kOffF2B", "kOffF3W", "kOffF3B"]
    assert [vals[n] for n in names] == list(L.PARAM_OFFSETS.values())
    assert vals["kFlatNumel"] == L.FLAT_NUMEL and vals["kParamNumel"] == L.PARAM_NUMEL == 62006
    hdr_c = (ROOT / "include" / "dnnb200.h").read_text()
    assert f"#define DNNB200_FLAT_NUMEL {L.FLAT_NUMEL}" in hdr_c


def test_state_dict_contract():
    m = Network()
    sd = m.state_dict()
    assert list(sd.keys()) == list(L.PARAM_SHAPES.keys())
    for k, shape in L.PARAM_SHAPES.items():
        assert tuple(sd[k].shape) == shape and sd[k].dtype == torch.float32
    assert sum(v.numel() for v in sd.values()) == 62006
    assert len(list(m.parameters())) == 10


def test_parameters_alias_flat_buffer_and_survive_moves():
    m = Network()
    assert m.conv1.weight.data_ptr() == m.flat_params.data_ptr()
    off = L.PARAM_OFFSETS["fc1.weight"]
    m.flat_params[off] = 123.0
    assert m.fc1.weight[0, 0].item() == 123.0
    m2 = copy.deepcopy(m).to("cpu")
    m2.flat_params[off] = 7.0
    assert m2.fc1.weight[0, 0].item() == 7.0 and m.fc1.weight[0, 0].item() == 123.0
    # gradients accumulate into the flat gradient buffer
    m.zero_grad()
    m(torch.randn(2, 3, 32, 32)).sum().backward()
    assert m.flat_grads.abs().sum() > 0
    assert m.fc3.bias.grad.data_ptr() == m.flat_grads[L.PARAM_OFFSETS["fc3.bias"]:].data_ptr()
    # padding never receives gradient
    assert m.flat_grads[~L.real_mask()].abs().sum() == 0


def test_pack_unpack_roundtrip():
    m = Network()
    flat = L.pack(m.state_dict())
    assert torch.equal(flat, m.flat_params)
    sd = L.unpack(flat)
    m2 = Network()
    m2.load_state_dict(sd)
    assert torch.equal(m2.flat_params, m.flat_params)


def test_default_init_matches_pytorch_bounds():
    torch.manual_seed(0)
    m = Network()
    for mod, fan_in in ((m.conv1, 75), (m.conv2, 150), (m.fc1, 400), (m.fc2, 120), (m.fc3, 84)):
        bound = fan_in ** -0.5
        assert mod.weight.abs().max() <= bound and mod.weight.abs().max() > 0.8 * bound
        assert mod.bias.abs().max() <= bound


def test_forward_matches_reference_network(reference_network_cls):
    torch.manual_seed(1)
    ref = reference_network_cls()
    ours = Network()
    ours.load_state_dict(ref.state_dict())        # reference checkpoint loads into ours ...
    ref.load_state_dict(ours.state_dict())        # ... and ours into the reference
    x = torch.randn(5, 3, 32, 32)
    assert torch.allclose(ours(x), ref(x), atol=1e-6)
    # and the gradients agree tensor by tensor
    y = torch.randint(0, 10, (5,))
    torch.nn.functional.cross_entropy(ref(x), y).backward()
    ours.zero_grad()
    torch.nn.functional.cross_entropy(ours(x), y).backward()
    for (n1, p1), (n2, p2) in zip(ref.named_parameters(), ours.named_parameters()):
        assert n1 == n2 and torch.allclose(p1.grad, p2.grad, atol=1e-6)


def test_top_level_import_path():
    import importlib
    import sys
    sys.path.insert(0, str(ROOT))
    mod = importlib.import_module("models.model")
    assert mod.Network is Network


def test_header_and_library_export_the_same_c_abi():
    """include/dnnb200.h is the contract: every function it declares is exported by the built library, and the
    library exports no dnnb200_* symbol the header does not declare."""
    import re
    import subprocess
    from pathlib import Path

    from distributed_neural_network_b200.ops import native

    root = Path(__file__).resolve().parent.parent
    lib_path = native.lib_path()
    if not Path(lib_path).exists():
        pytest.skip("native library not built")
    header = (root / "include" / "dnnb200.h").read_text()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(dnnb200_[a-z0-9_]+)\s*\(", header))
    out = subprocess.run(["nm", "-D", "--defined-only", str(lib_path)], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.split() and ln.split()[-1].startswith("dnnb200_")}
    assert declared, "no declarations parsed from the header"
    assert declared - exported == set(), f"declared but not exported: {sorted(declared - exported)}"
    assert exported - declared == set(), f"exported but not declared: {sorted(exported - declared)}"


def test_python_mirror_of_the_control_block_matches_the_header():
    import ctypes
    import re
    from pathlib import Path

    from distributed_neural_network_b200.ops import native

    header = (Path(__file__).resolve().parent.parent / "include" / "dnnb200.h").read_text()
    body = re.search(r"typedef struct dnnb200_ctl \{(.*?)\} dnnb200_ctl;", header, flags=re.S).group(1)
    ctype = {"unsigned long long": ctypes.c_ulonglong, "float": ctypes.c_float, "unsigned int": ctypes.c_uint}
    want = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        m = re.match(r"(unsigned long long|unsigned int|float)\s+(.*)", decl)
        assert m, decl
        for name in m.group(2).split(","):
            want.append((name.strip(), ctype[m.group(1)]))
    assert [(n, t) for n, t in native.Ctl._fields_] == want
    assert ctypes.sizeof(native.Ctl) == 56


def test_header_is_plain_c_and_links_from_a_c_program(tmp_path):
    """include/dnnb200.h must be usable from C (the ABI is extern "C", integer status codes): compile a C99
    program with -pedantic against it, link the in-tree library, and run the one entry point that needs no GPU."""
    import shutil
    import subprocess
    from pathlib import Path

    from distributed_neural_network_b200.ops import flat, native

    root = Path(__file__).resolve().parent.parent
    lib_dir = Path(native.lib_path()).parent
    if shutil.which("gcc") is None or not Path(native.lib_path()).exists():
        pytest.skip("needs gcc and the built library")
    src = tmp_path / "abi_check.c"
    src.write_text(
        '#include <stdint.h>\n#include <stdio.h>\n#include "dnnb200.h"\n'
        "int main(void) {\n"
        "  int offs[10], numel, slice, nslices;\n"
        "  if (dnnb200_layout(offs, &numel, &slice, &nslices) != 0) return 1;\n"
        '  printf("%d %d %d %d %d\\n", numel, slice, nslices, offs[9], dnnb200_version());\n'
        "  return 0;\n}\n")
    exe = tmp_path / "abi_check"
    cc = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", f"-I{root / 'include'}", str(src),
                         f"-L{lib_dir}", "-ldnnb200", f"-Wl,-rpath,{lib_dir}", "-o", str(exe)],
                        capture_output=True, text=True)
    assert cc.returncode == 0, cc.stderr
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0, out.stderr
    numel, slice_, nslices, last_off, version = map(int, out.stdout.split())
    assert numel == flat.FLAT_NUMEL and last_off == list(flat.PARAM_OFFSETS.values())[-1]
    assert slice_ * nslices >= numel and version >= 100
