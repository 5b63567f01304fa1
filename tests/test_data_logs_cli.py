"""Host-side parity: partitioning, log files, CLI, checkpoint (SURVEY §4 unit/host level; W5-W7, W10, W11)."""
import ast
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

from distributed_neural_network_b200 import Network
from distributed_neural_network_b200.ops import flat as L
from distributed_neural_network_b200.utils import checkpoint as ckpt
from distributed_neural_network_b200.utils import data as D
from distributed_neural_network_b200.utils import logs

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")


def _reference_function(name):
    src_file = REF / "data_parallelism_train.py"
    if not src_file.exists():
        pytest.skip("reference checkout not available")
    tree = ast.parse(src_file.read_text())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == name)
    ns = {}
    import torch.utils.data as data
    ns["data"] = data
    exec(compile(ast.Module([fn], []), str(src_file), "exec"), ns)
    return ns[name]


@pytest.mark.parametrize("n,size", [(50000, 4), (50000, 8), (50000, 3), (1001, 4), (17, 5)])
def test_partition_dataset_index_exact(n, size):
    ref_fn = _reference_function("partition_dataset")
    ds = list(range(n))
    for rank in range(1, size):
        assert D.partition_dataset(ds, rank, size).indices == list(ref_fn(ds, rank, size).indices)


def test_partition_drops_remainder_and_is_contiguous():
    ds = D.synthetic(103)
    parts = [D.partition_dataset(ds, r, 4) for r in (1, 2, 3)]
    assert [len(p) for p in parts] == [34, 34, 34]
    assert parts[1].indices[0] == 34 and parts[2].indices[-1] == 101
    m = parts[1].materialize()
    assert torch.equal(m.images, ds.images[34:68])


def test_transform_matches_torchvision_semantics():
    import numpy as np
    u8 = np.random.RandomState(0).randint(0, 256, (4, 32, 32, 3), dtype=np.uint8)
    t = D.transform(u8)
    assert t.shape == (4, 3, 32, 32) and t.dtype == torch.float32
    assert t.min() >= -1 and t.max() <= 1
    want = (torch.from_numpy(u8).permute(0, 3, 1, 2).float() / 255 - 0.5) / 0.5
    assert torch.allclose(t, want, atol=1e-6)


def test_synthetic_is_deterministic_and_shaped():
    a, b = D.synthetic(64, seed=3), D.synthetic(64, seed=3)
    assert torch.equal(a.images, b.images) and torch.equal(a.labels, b.labels)
    assert a.images.shape == (64, 3, 32, 32) and a.labels.dtype == torch.int64 and int(a.labels.max()) <= 9


def test_epoch_tensor_covers_every_sample_once():
    ds = D.synthetic(37)
    ds.labels = torch.arange(37)
    x, y, tail = D.epoch_tensor(ds, 8, shuffle=True, generator=torch.Generator().manual_seed(0))
    assert x.shape == (4, 8, 3, 32, 32) and tail[1].numel() == 5
    assert sorted(torch.cat([y.flatten(), tail[1]]).tolist()) == list(range(37))


def test_log_files_have_reference_names_and_labels(tmp_path):
    t = logs.Timers(1.5, 2.5, 3.5, 4.5, 5.5)
    p = logs.write_parent_log(t, 16, 25, 4, str(tmp_path))
    c = logs.write_children_log(t, 16, 25, 4, str(tmp_path))
    assert os.path.basename(p) == "bs16_log_epochs25_proc4_parent.txt"
    assert os.path.basename(c) == "bs16_log_epochs25_proc4_children.txt"
    pl, cl = open(p).read().splitlines(), open(c).read().splitlines()
    assert [l.split(":")[0] for l in pl] == ["Eval data loading time", "Time spent on evaluation",
                                              "Time spent on parent communication and param sync"]
    assert [l.split(":")[0] for l in cl] == ["Train data loading time", "Time spent on training",
                                              "Time spent on children communication"]
    assert float(pl[1].split(": ")[1]) == 3.5 and float(cl[1].split(": ")[1]) == 2.5
    ref_p = REF / "log" / "bs16_log_epochs25_proc4_parent.txt"
    if ref_p.exists():  # byte-compatible labels with the committed reference logs
        assert [l.split(":")[0] for l in ref_p.read_text().splitlines()] == [l.split(":")[0] for l in pl]
        ref_c = REF / "log" / "bs16_log_epochs25_proc4_children.txt"
        assert [l.split(":")[0] for l in ref_c.read_text().splitlines()] == [l.split(":")[0] for l in cl]


def test_null_run_has_neptune_surface_and_no_token_in_repo():
    run = logs.init_run(enabled=False)
    run["parameters"] = {"lr": 0.001}
    run["train/loss"].append(1.25)
    run["val/acc"].append(50)
    run.stop()
    assert run.series["train/loss"] == [1.25] and run.values["parameters"]["lr"] == 0.001
    for f in list(ROOT.glob("*.py")) + list((ROOT / "distributed_neural_network_b200").rglob("*.py")):
        assert "api_token=\"ey" not in f.read_text()


def test_cli_flags_match_reference_and_are_typed():
    out = subprocess.run([sys.executable, str(ROOT / "data_parallelism_train.py"), "--help"], capture_output=True,
                         text=True, cwd=ROOT)
    assert out.returncode == 0
    for flag in ("--lr", "--momentum", "--batch-size", "--epochs", "--nb-proc", "--failure-probability",
                 "--failure-duration"):
        assert flag in out.stdout
    import argparse
    from distributed_neural_network_b200.utils import cli
    p = argparse.ArgumentParser()
    cli.add_reference_flags(p, 25, True)
    cli.add_new_flags(p, "step-allreduce")
    a = p.parse_args([])
    assert (a.lr, a.momentum, a.bs, a.epochs, a.nb_proc, a.failure_probability, a.failure_duration) == \
        (0.001, 0.9, 16, 25, 4, 0.0, 0.0)
    a = p.parse_args(["--lr", "0.1", "--batch-size", "8"])
    assert isinstance(a.lr, float) and isinstance(a.bs, int)       # the reference passes strings (Q5)
    p2 = argparse.ArgumentParser()
    cli.add_reference_flags(p2, 10, False)
    assert p2.parse_args([]).epochs == 10                           # model_replication_train.py default


def test_checkpoint_is_reference_format_and_resumes_exactly(tmp_path, reference_network_cls):
    torch.manual_seed(0)
    m = Network()
    mom = torch.randn(L.FLAT_NUMEL)
    path = str(tmp_path / "ck.pt")
    ckpt.save(path, m, mom, epoch=3)
    sd = torch.load(path, weights_only=True)
    assert list(sd.keys()) == list(L.PARAM_SHAPES.keys())
    ref = reference_network_cls()
    ref.load_state_dict(sd)                                          # loads into the reference class unchanged
    m2, mom2 = Network(), torch.zeros(L.FLAT_NUMEL)
    meta = ckpt.load(path, m2, mom2)
    assert meta["epoch"] == 3 and torch.equal(m2.flat_params, m.flat_params) and torch.equal(mom2, mom)


def test_sweep_report_reproduces_the_reference_tables(tmp_path):
    """utils/report.py on timing files with the reference's committed values (log_epochs25_proc{3,8}_*, bs{1,64}_...proc4,
    SURVEY section 6.2): same derived throughput, the un-prefixed and the bs16_ run of 4 processes kept apart."""
    from distributed_neural_network_b200.utils import report

    def put(name, a, b, c, labels):
        (tmp_path / name).write_text("".join(f"{lab}: {v}\n" for lab, v in zip(labels, (a, b, c))))

    ch = ("Train data loading time", "Time spent on training", "Time spent on children communication")
    pa = ("Eval data loading time", "Time spent on evaluation", "Time spent on parent communication and param sync")
    put("log_epochs25_proc3_children.txt", 4.27, 316.52, 0.161, ch)
    put("log_epochs25_proc3_parent.txt", 1.94, 56.45, 0.179, pa)
    put("log_epochs25_proc8_children.txt", 18.52, 1566.32, 0.694, ch)
    put("log_epochs25_proc8_parent.txt", 7.75, 58.00, 0.778, pa)
    put("log_epochs25_proc4_children.txt", 6.19, 734.60, 0.235, ch)
    put("log_epochs25_proc4_parent.txt", 2.57, 56.74, 0.255, pa)
    put("bs16_log_epochs25_proc4_children.txt", 6.31, 701.82, 0.240, ch)
    put("bs16_log_epochs25_proc4_parent.txt", 2.54, 56.69, 0.227, pa)
    put("bs1_log_epochs25_proc4_children.txt", 6.07, 1167.29, 0.166, ch)
    put("bs1_log_epochs25_proc4_parent.txt", 2.57, 158.62, 0.282, pa)
    (tmp_path / "notes.txt").write_text("not a timing file\n")
    rows = report.read_logs(str(tmp_path))
    assert len(rows) == 5
    for r in rows.values():
        report.derive(r, 50000, "reference")
    t1, t2 = report.tables(rows)
    assert [r.nb_proc for r in t1] == [3, 4, 8] and not t1[1].bs_in_name
    assert round(t1[0].img_per_s_aggregate) == 3949 and round(t1[2].img_per_s_aggregate) == 798
    assert round(t1[0].comm_ms_per_epoch_child, 1) == 6.4 and round(t1[2].comm_ms_per_epoch_parent, 1) == 31.1
    assert abs(t1[0].total_s - (316.52 + 56.45 + 4.27 + 1.94)) < 1e-9
    assert list(t2) == [4] and [r.batch_size for r in t2[4]] == [1, 16] and t2[4][1].bs_in_name
    assert round(t2[4][0].img_per_s_aggregate) == 1071 and round(t2[4][1].img_per_s_aggregate) == 1781
    sym = report.derive(report.RunRow(nb_proc=8, batch_size=16, epochs=5, train_s=10.0), 50000, "symmetric")
    assert sym.images_per_trainer_epoch == 6250 and round(sym.img_per_s_aggregate) == 25000
    text = report.render(str(tmp_path))
    assert "Table 1" in text and "Table 2" in text and "| 3 | 316.52 | 56.45 |" in text
    assert report.main([str(tmp_path), "--json"]) == 0


def test_report_writes_accuracy_and_loss_curves_as_svg(tmp_path):
    """The report's Fig. 1 / 2 from the per-epoch JSONL, as a stand-alone SVG (no plotting package on the box)."""
    import json
    import xml.dom.minidom

    from distributed_neural_network_b200.utils import report
    for name, shift in (("dp4_bs16", 0.0), ("dp8_bs16", 3.0)):
        with open(tmp_path / f"{name}.jsonl", "w") as f:
            for e in range(6):
                f.write(json.dumps({"epoch": e, "train_loss": 2.3 - 0.2 * e, "val_loss": 2.2 - 0.18 * e,
                                    "val_acc": 10 + 9 * e - shift}) + "\n")
    curves = report.read_curves(sorted(str(p) for p in tmp_path.glob("*.jsonl")))
    assert list(curves) == ["dp4_bs16", "dp8_bs16"] and len(curves["dp4_bs16"]) == 6
    out = tmp_path / "curves.svg"
    assert report.main([str(tmp_path), "--svg", str(out)]) == 0
    doc = xml.dom.minidom.parse(str(out))
    assert len(doc.getElementsByTagName("polyline")) == 2 + 4        # accuracy per run; val + train loss per run
    assert "validation accuracy" in out.read_text()


def test_cifar_readers_agree_on_pickled_binary_and_tarball_layouts(tmp_path):
    """The three on-disk forms a user may have (extracted python batches as the reference's torchvision download leaves
    them, the dataset's binary version, the undecompressed tarball) must give the same tensors, in [-1, 1], NCHW."""
    import pickle
    import tarfile

    import numpy as np

    rng = np.random.default_rng(0)
    per = 4
    data = [rng.integers(0, 256, size=(per, 3072), dtype=np.uint8) for _ in range(6)]      # 5 train files + test
    labels = [rng.integers(0, 10, size=per).tolist() for _ in range(6)]
    names = [f"data_batch_{i}" for i in range(1, 6)] + ["test_batch"]

    py_root, bin_root, tar_root = tmp_path / "py", tmp_path / "bin", tmp_path / "tar"
    (py_root / D.CIFAR_DIR).mkdir(parents=True)
    (bin_root / D.CIFAR_BIN_DIR).mkdir(parents=True)
    tar_root.mkdir()
    for n, x, y in zip(names, data, labels):
        with open(py_root / D.CIFAR_DIR / n, "wb") as f:
            pickle.dump({"data": x, "labels": y}, f)
        rec = np.concatenate([np.asarray(y, dtype=np.uint8)[:, None], x], axis=1)
        rec.tofile(bin_root / D.CIFAR_BIN_DIR / f"{n}.bin")
    with tarfile.open(tar_root / D.CIFAR_TARBALL, "w:gz") as tar:
        tar.add(py_root / D.CIFAR_DIR, arcname=D.CIFAR_DIR)

    for train, n_img in ((True, 5 * per), (False, per)):
        got = [D.cifar10(str(r), train=train, allow_synthetic=False) for r in (py_root, bin_root, tar_root)]
        for ds in got:
            assert ds.source == "cifar10" and ds.images.shape == (n_img, 3, 32, 32) and ds.images.dtype == torch.float32
            assert ds.labels.dtype == torch.int64 and float(ds.images.min()) >= -1.0 and float(ds.images.max()) <= 1.0
        assert torch.equal(got[0].images, got[1].images) and torch.equal(got[0].images, got[2].images)
        assert torch.equal(got[0].labels, got[1].labels) and torch.equal(got[0].labels, got[2].labels)
    want = torch.from_numpy(data[0][0].reshape(3, 32, 32).astype(np.float32)) / 255.0 * 2 - 1
    assert torch.allclose(D.cifar10(str(bin_root), train=True, allow_synthetic=False).images[0], want, atol=1e-6)
    (bin_root / D.CIFAR_BIN_DIR / "test_batch.bin").write_bytes(b"\\0" * 100)
    with pytest.raises(ValueError, match="3073-byte record"):
        D.cifar10(str(bin_root), train=False, allow_synthetic=False)
    with pytest.raises(FileNotFoundError):
        D.cifar10(str(tmp_path / "nothing"), train=True, allow_synthetic=False)
