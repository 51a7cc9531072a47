"""Convergence curves from experiment JSON files written by the examples.

Counterpart of reference ``examples/utils/convergence_plots.py:11-211`` (``plot_rounds_convergence``,
``plot_brainage_convergence``).  The series extraction is independent of any plotting package; figures are drawn
with matplotlib when it is importable and otherwise the same series are written as CSV next to the requested
output (this image ships no matplotlib).
"""
import csv
import datetime
import json
import os
import sys


def convert_to_timestamp(t):
    """RFC-3339 string (as produced by ``MessageToDict`` for ``google.protobuf.Timestamp``) -> POSIX seconds."""
    if isinstance(t, (int, float)):
        return float(t)
    t = t.rstrip("Z")
    if "." in t:
        head, frac = t.split(".")
        t = head + "." + (frac + "000000")[:6]
        fmt = "%Y-%m-%dT%H:%M:%S.%f"
    else:
        fmt = "%Y-%m-%dT%H:%M:%S"
    return datetime.datetime.strptime(t, fmt).replace(tzinfo=datetime.timezone.utc).timestamp()


def community_series(experiment, metric="accuracy", split="test_evaluation"):
    """[(global_iteration, mean metric over learners, wall-clock seconds since round 1 started)]."""
    started = {}
    for md in experiment.get("federation_runtime_metadata", {}).get("metadata", []):
        if "completed_at" in md:
            started[int(md.get("global_iteration", 0))] = (convert_to_timestamp(md["started_at"]),
                                                           convert_to_timestamp(md["completed_at"]))
    t0 = min((s for s, _ in started.values()), default=0.0)
    out = []
    for ev in experiment.get("community_model_results", {}).get("community_evaluation", []):
        it = int(ev.get("global_iteration", 0))
        vals = [float(e[split]["metric_values"][metric]) for e in ev.get("evaluations", {}).values()
                if metric in e.get(split, {}).get("metric_values", {})]
        if vals:
            done = started.get(it, (t0, t0))[1]
            out.append((it, sum(vals) / len(vals), max(0.0, done - t0)))
    return sorted(out)


def round_durations(experiment):
    """[(global_iteration, seconds, aggregation milliseconds)] from the controller's runtime metadata."""
    rows = []
    for md in experiment.get("federation_runtime_metadata", {}).get("metadata", []):
        if "completed_at" not in md:
            continue
        secs = convert_to_timestamp(md["completed_at"]) - convert_to_timestamp(md["started_at"])
        rows.append((int(md.get("global_iteration", 0)), secs, float(md.get("model_aggregation_total_duration_ms", 0.0))))
    return sorted(rows)


def _load(files):
    if isinstance(files, (str, os.PathLike)):
        files = [files]
    return [(os.path.basename(str(f)), json.load(open(f))) for f in files]


def _emit(series_by_label, xlabel, ylabel, out):
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        path = os.path.splitext(out)[0] + ".csv"
        with open(path, "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["experiment", xlabel, ylabel])
            for label, pts in series_by_label.items():
                w.writerows([label, x, y] for x, y in pts)
        return path
    fig, ax = plt.subplots(figsize=(7, 4))
    for label, pts in series_by_label.items():
        ax.plot([p[0] for p in pts], [p[1] for p in pts], marker="o", label=label)
    ax.set_xlabel(xlabel); ax.set_ylabel(ylabel); ax.grid(True, alpha=0.3); ax.legend()
    fig.tight_layout(); fig.savefig(out, dpi=150)
    return out


def plot_rounds_convergence(files, metric="accuracy", out="/tmp/metis/rounds_convergence.png"):
    series = {name: [(it, v) for it, v, _ in community_series(exp, metric)] for name, exp in _load(files)}
    return _emit(series, "federation round", metric, out)


def plot_brainage_convergence(files, metric="mae", show_global_models_exchanged=False, show_processing_time=False,
                              out="/tmp/metis/brainage_convergence.png"):
    """Regression convergence against rounds, wall-clock time, or models exchanged (2 x learners per round)."""
    series = {}
    for name, exp in _load(files):
        n_learners = max(1, len(exp.get("learners_descriptor", {}).get("learner", [])))
        pts = community_series(exp, metric)
        if show_processing_time:
            series[name] = [(secs, v) for _, v, secs in pts]
        elif show_global_models_exchanged:
            series[name] = [(it * 2 * n_learners, v) for it, v, _ in pts]
        else:
            series[name] = [(it, v) for it, v, _ in pts]
    xlabel = "seconds" if show_processing_time else ("models exchanged" if show_global_models_exchanged else "federation round")
    return _emit(series, xlabel, metric, out)


if __name__ == "__main__":
    print(plot_rounds_convergence(sys.argv[1:] or ["/tmp/metis/experiment.json"]))
