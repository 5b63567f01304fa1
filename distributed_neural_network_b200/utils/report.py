"""Sweep report: the timing files of a log directory as the two tables of the reference's project report.

The reference publishes a process-count sweep (Table 1) and a batch-size sweep (Table 2) that were assembled by hand
from the files ``run_training.sh`` leaves under ``log/`` (three cumulative timers per role and run, reference
data_parallelism_train.py:103-104,126-129,143-152). This module does that assembly: it reads every
``[bs{B}_]log_epochs{E}_proc{P}_{parent|children}.txt`` of a directory -- this framework's and the reference's
committed ones alike, the file format is the same -- adds the derived throughput columns and, when
a per-epoch JSONL of this framework lies next to them, accuracy, loss and the device-timed step and exchange figures.

    python -m distributed_neural_network_b200.utils.report log/            # markdown on stdout
    python -m distributed_neural_network_b200.utils.report log/ --json     # the rows as JSON
    python -m distributed_neural_network_b200.utils.report log/ --svg curves.svg   # + accuracy / loss curves (Fig. 1 / 2)

Shard sizes follow the topology that wrote the files: ``--topology reference`` (rank 0 does not train, P-1 shards of
``train_size // (P-1)`` images, reference :49-53) or ``symmetric`` (all P ranks train ``train_size // P``; the default
of this framework's step-allreduce mode).
"""
from __future__ import annotations

import argparse
import json
import os
import re
import sys
from dataclasses import asdict, dataclass
from typing import Dict, List, Optional, Tuple

Key = Tuple[int, int, int, bool]

_NAME = re.compile(r"^(?:bs(?P<bs>\d+)_)?log_epochs(?P<epochs>\d+)_proc(?P<proc>\d+)_(?P<role>parent|children)\.txt$")
_DEFAULT_BS = 16      # files without a bs prefix were written with the default --batch-size (reference :263)


@dataclass
class RunRow:
    nb_proc: int
    batch_size: int
    epochs: int
    bs_in_name: bool = True                    # False: the file name carries no bs prefix (the reference's process sweep)
    train_s: Optional[float] = None            # children: "Time spent on training"
    child_comm_s: Optional[float] = None       # children: "... children communication"
    child_load_s: Optional[float] = None       # children: "Train data loading time"
    eval_s: Optional[float] = None             # parent: "Time spent on evaluation"
    parent_comm_s: Optional[float] = None      # parent: "... parent communication and param sync"
    parent_load_s: Optional[float] = None      # parent: "Eval data loading time"
    images_per_trainer_epoch: Optional[int] = None
    img_per_s_per_trainer: Optional[float] = None
    img_per_s_aggregate: Optional[float] = None
    comm_ms_per_epoch_child: Optional[float] = None
    comm_ms_per_epoch_parent: Optional[float] = None
    total_s: Optional[float] = None            # the report's "Training Time": training + evaluation + both loads
    val_acc: Optional[float] = None            # from a JSONL of this framework, last epoch
    val_loss: Optional[float] = None
    step_us: Optional[float] = None
    allreduce_us: Optional[float] = None


def _three_floats(path: str) -> Tuple[float, float, float]:
    vals = []
    with open(path) as f:
        for line in f:
            if ":" in line:
                vals.append(float(line.rsplit(":", 1)[1]))
    if len(vals) != 3:
        raise ValueError(f"{path}: expected three 'label: seconds' lines, found {len(vals)}")
    return vals[0], vals[1], vals[2]


def read_logs(log_dir: str) -> Dict[Key, RunRow]:
    """{(nb_proc, batch_size, epochs, bs in the file name): RunRow} with the six timers the files of ``log_dir`` hold.
    (The reference's directory has ``log_epochs25_proc4_*`` AND ``bs16_log_epochs25_proc4_*``: two runs, kept apart.)"""
    rows: Dict[Key, RunRow] = {}
    for name in sorted(os.listdir(log_dir)):
        m = _NAME.match(name)
        if not m:
            continue
        bs = int(m["bs"]) if m["bs"] else _DEFAULT_BS
        key = (int(m["proc"]), bs, int(m["epochs"]), bool(m["bs"]))
        row = rows.setdefault(key, RunRow(nb_proc=key[0], batch_size=bs, epochs=key[2], bs_in_name=key[3]))
        a, b, c = _three_floats(os.path.join(log_dir, name))
        if m["role"] == "children":
            row.child_load_s, row.train_s, row.child_comm_s = a, b, c
        else:
            row.parent_load_s, row.eval_s, row.parent_comm_s = a, b, c
    return rows


def derive(row: RunRow, train_size: int = 50000, topology: str = "reference") -> RunRow:
    trainers = row.nb_proc - 1 if topology == "reference" else row.nb_proc
    if trainers < 1:
        trainers = 1
    row.images_per_trainer_epoch = train_size // trainers
    if row.train_s:
        row.img_per_s_per_trainer = row.epochs * row.images_per_trainer_epoch / row.train_s
        row.img_per_s_aggregate = row.img_per_s_per_trainer * trainers
    if row.child_comm_s is not None:
        row.comm_ms_per_epoch_child = 1e3 * row.child_comm_s / row.epochs
    if row.parent_comm_s is not None:
        row.comm_ms_per_epoch_parent = 1e3 * row.parent_comm_s / row.epochs
    parts = [row.train_s, row.eval_s, row.child_load_s, row.parent_load_s]
    if all(p is not None for p in parts):
        row.total_s = float(sum(parts))
    return row


def attach_jsonl(rows: Dict[Key, RunRow], log_dir: str) -> None:
    """``dp{P}_bs{B}.jsonl`` (or any ``*.jsonl`` whose records carry nothing but one run) -> last epoch's figures."""
    for name in sorted(os.listdir(log_dir)):
        if not name.endswith(".jsonl"):
            continue
        m = re.match(r"^dp(\d+)_bs(\d+)\.jsonl$", name)
        recs = []
        with open(os.path.join(log_dir, name)) as f:
            for line in f:
                line = line.strip()
                if line:
                    recs.append(json.loads(line))
        if not recs or not m:
            continue
        last = recs[-1]
        for key, row in rows.items():
            if key[0] == int(m[1]) and key[1] == int(m[2]):
                row.val_acc, row.val_loss = last.get("val_acc"), last.get("val_loss")
                row.step_us, row.allreduce_us = last.get("step_us"), last.get("allreduce_us")


def _fmt(v, spec: str) -> str:
    return "" if v is None else format(v, spec)


def _table(title: str, rows: List[RunRow], first: str) -> str:
    out = [f"### {title}", "",
           f"| {first} | train s | eval s | total s | img/s per trainer | img/s aggregate | comm ms/epoch child / parent |"
           " val acc % | val loss | step us | exchange us |",
           "|--:|--:|--:|--:|--:|--:|---|--:|--:|--:|--:|"]
    for r in rows:
        lead = r.nb_proc if first == "processes" else r.batch_size
        out.append(f"| {lead} | {_fmt(r.train_s, '.2f')} | {_fmt(r.eval_s, '.2f')} | {_fmt(r.total_s, '.1f')} | "
                   f"{_fmt(r.img_per_s_per_trainer, '.0f')} | {_fmt(r.img_per_s_aggregate, '.0f')} | "
                   f"{_fmt(r.comm_ms_per_epoch_child, '.1f')} / {_fmt(r.comm_ms_per_epoch_parent, '.1f')} | "
                   f"{_fmt(r.val_acc, '.2f')} | {_fmt(r.val_loss, '.3f')} | {_fmt(r.step_us, '.1f')} | "
                   f"{_fmt(r.allreduce_us, '.2f')} |")
    return "\n".join(out)


def tables(rows: Dict[Key, RunRow]) -> Tuple[List[RunRow], Dict[int, List[RunRow]]]:
    """(process sweep at the default batch size -- the report's Table 1 --, {nb_proc: batch sweep} -- Table 2).
    Where a run exists with and without the bs prefix, Table 1 takes the un-prefixed file and Table 2 the prefixed."""
    table1: Dict[int, RunRow] = {}
    for r in sorted(rows.values(), key=lambda r: (r.nb_proc, r.bs_in_name, -r.epochs)):
        if r.batch_size == _DEFAULT_BS:
            table1.setdefault(r.nb_proc, r)
    sweeps: Dict[int, Dict[int, RunRow]] = {}
    for r in sorted(rows.values(), key=lambda r: (r.nb_proc, r.batch_size, not r.bs_in_name, -r.epochs)):
        sweeps.setdefault(r.nb_proc, {}).setdefault(r.batch_size, r)
    table2 = {p: [rs[b] for b in sorted(rs)] for p, rs in sweeps.items() if len(rs) > 1}
    return [table1[p] for p in sorted(table1)], table2


def render(log_dir: str, train_size: int = 50000, topology: str = "reference") -> str:
    rows = read_logs(log_dir)
    for r in rows.values():
        derive(r, train_size, topology)
    attach_jsonl(rows, log_dir)
    table1, table2 = tables(rows)
    parts = [f"## Sweep report for `{log_dir}` ({len(rows)} runs, {topology} topology, {train_size} training images)", ""]
    if table1:
        parts += [_table(f"Process sweep at batch size {_DEFAULT_BS} (reference report: Table 1)", table1, "processes"), ""]
    for p, rs in table2.items():
        parts += [_table(f"Batch-size sweep at {p} processes (reference report: Table 2)", rs, "batch size"), ""]
    if table1:
        best = max((r for r in rows.values() if r.img_per_s_aggregate), key=lambda r: r.img_per_s_aggregate, default=None)
        if best is not None:
            parts.append(f"Best aggregate throughput: {best.img_per_s_aggregate:.0f} img/s at {best.nb_proc} processes, "
                         f"batch size {best.batch_size}.")
    return "\n".join(parts)


# ---------------------------------------------------------------------------
# Figures 1 / 2 of the reference's report: validation accuracy and loss per epoch. No plotting package on the box,
# so the curves are written as a small stand-alone SVG.
# ---------------------------------------------------------------------------
_COLOURS = ("#1f77b4", "#d62728", "#2ca02c", "#9467bd", "#ff7f0e", "#8c564b", "#17becf", "#7f7f7f")


def read_curves(paths: List[str]) -> Dict[str, List[dict]]:
    """{label: per-epoch records} of the given JSONL files (label = file name without the extension)."""
    curves: Dict[str, List[dict]] = {}
    for path in paths:
        recs = []
        with open(path) as f:
            for line in f:
                line = line.strip()
                if line:
                    rec = json.loads(line)
                    if "epoch" in rec:
                        recs.append(rec)
        if recs:
            curves[os.path.splitext(os.path.basename(path))[0]] = sorted(recs, key=lambda r: r["epoch"])
    return curves


def _panel(x0: float, y0: float, w: float, h: float, title: str, series: Dict[str, List[Tuple[float, float]]]) -> List[str]:
    pts = [p for s in series.values() for p in s]
    if not pts:
        return []
    xmin, xmax = min(p[0] for p in pts), max(p[0] for p in pts)
    ymin, ymax = min(p[1] for p in pts), max(p[1] for p in pts)
    if xmax == xmin:
        xmax = xmin + 1
    if ymax == ymin:
        ymax = ymin + 1
    pad = 0.05 * (ymax - ymin)
    ymin, ymax = ymin - pad, ymax + pad
    sx = lambda x: x0 + (x - xmin) / (xmax - xmin) * w          # noqa: E731
    sy = lambda y: y0 + h - (y - ymin) / (ymax - ymin) * h      # noqa: E731
    out = [f'<rect x="{x0:.1f}" y="{y0:.1f}" width="{w:.1f}" height="{h:.1f}" fill="none" stroke="#444"/>',
           f'<text x="{x0 + w / 2:.1f}" y="{y0 - 8:.1f}" text-anchor="middle" font-size="13">{title}</text>']
    for k in range(5):
        yv = ymin + (ymax - ymin) * k / 4
        out.append(f'<line x1="{x0:.1f}" y1="{sy(yv):.1f}" x2="{x0 + w:.1f}" y2="{sy(yv):.1f}" stroke="#ddd"/>')
        out.append(f'<text x="{x0 - 6:.1f}" y="{sy(yv) + 4:.1f}" text-anchor="end" font-size="10">{yv:.3g}</text>')
    for xv in sorted({int(round(xmin + (xmax - xmin) * k / 4)) for k in range(5)}):
        out.append(f'<text x="{sx(xv):.1f}" y="{y0 + h + 14:.1f}" text-anchor="middle" font-size="10">{xv}</text>')
    out.append(f'<text x="{x0 + w / 2:.1f}" y="{y0 + h + 30:.1f}" text-anchor="middle" font-size="11">epoch</text>')
    for i, (label, s) in enumerate(series.items()):
        colour = _COLOURS[i % len(_COLOURS)]
        path = " ".join(f"{sx(x):.1f},{sy(y):.1f}" for x, y in s)
        out.append(f'<polyline points="{path}" fill="none" stroke="{colour}" stroke-width="1.8"/>')
        out.append(f'<text x="{x0 + 8:.1f}" y="{y0 + 16 + 14 * i:.1f}" font-size="11" fill="{colour}">{label}</text>')
    return out


def curves_svg(curves: Dict[str, List[dict]]) -> str:
    """Validation accuracy (left) and validation + training loss (right) against the epoch, one line per run."""
    acc = {k: [(r["epoch"], r["val_acc"]) for r in v if r.get("val_acc") is not None] for k, v in curves.items()}
    loss: Dict[str, List[Tuple[float, float]]] = {}
    for k, v in curves.items():
        loss[f"{k} val"] = [(r["epoch"], r["val_loss"]) for r in v if r.get("val_loss") is not None]
        loss[f"{k} train"] = [(r["epoch"], r["train_loss"]) for r in v if r.get("train_loss") is not None]
    body = _panel(60, 40, 380, 260, "validation accuracy (%)", {k: s for k, s in acc.items() if s})
    body += _panel(520, 40, 380, 260, "loss", {k: s for k, s in loss.items() if s})
    return ('<svg xmlns="http://www.w3.org/2000/svg" width="940" height="350" font-family="sans-serif">'
            '<rect width="100%" height="100%" fill="white"/>' + "".join(body) + "</svg>\n")


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("log_dir")
    ap.add_argument("--train-size", type=int, default=50000)
    ap.add_argument("--topology", choices=("reference", "symmetric"), default="reference")
    ap.add_argument("--json", action="store_true", help="print the rows as JSON instead of markdown")
    ap.add_argument("--svg", metavar="FILE", help="also write accuracy / loss curves of every *.jsonl in log_dir (the "
                                                  "report's Fig. 1 / 2) to FILE")
    a = ap.parse_args(argv)
    if a.svg:
        curves = read_curves([os.path.join(a.log_dir, n) for n in sorted(os.listdir(a.log_dir)) if n.endswith(".jsonl")])
        with open(a.svg, "w") as f:
            f.write(curves_svg(curves))
    if a.json:
        rows = read_logs(a.log_dir)
        for r in rows.values():
            derive(r, a.train_size, a.topology)
        attach_jsonl(rows, a.log_dir)
        print(json.dumps([asdict(r) for _, r in sorted(rows.items())], indent=1))
    else:
        print(render(a.log_dir, a.train_size, a.topology))
    return 0


if __name__ == "__main__":
    sys.exit(main())
