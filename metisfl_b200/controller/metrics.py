"""Prometheus endpoint of the controller (the reference has logs only).

``METISFL_PROMETHEUS_PORT=9464`` (or ``ControllerServicer(..., prometheus_port=...)``) starts an HTTP endpoint whose
values are read from the controller core at scrape time - no bookkeeping on the federation's hot path:

    metisfl_global_iteration                       current federation round
    metisfl_learners                               learners in the federation
    metisfl_fabric_learners                        of them, ranks of an on-box device fabric
    metisfl_model_bytes_received_total             model payload bytes that came in through MarkTaskCompleted
    metisfl_aggregation_failures_total             aggregations that closed a round without a new community model
    metisfl_last_aggregation_duration_ms           duration of the newest aggregation (sum of its blocks)
    metisfl_last_aggregation_device_timed          1 when that duration is a CUDA-event time from the fabric
    metisfl_last_round_duration_seconds            completed_at - started_at of the newest finished round
"""
from __future__ import annotations

import os
import threading

from metisfl_b200.utils.metis_logger import MetisLogger


class ControllerMetrics:
    def __init__(self, servicer, port: int = None, addr: str = "0.0.0.0"):
        self._servicer = servicer
        self._port = int(port if port is not None else os.environ.get("METISFL_PROMETHEUS_PORT", "0") or 0)
        self._addr = addr
        self._httpd = None
        self._thread = None
        self.registry = None

    @property
    def enabled(self) -> bool:
        return self._port > 0

    def snapshot(self) -> dict:
        """The exported values as a plain dict (also what the collector reads)."""
        s = self._servicer
        core = s.core
        out = {"global_iteration": float(core.global_iteration()), "learners": float(len(core.get_learners())),
               "fabric_learners": float(len(getattr(s, "_fabric_of", {}))),
               "model_bytes_received_total": float(s.controller_rx_model_bytes),
               "aggregation_failures_total": float(core.aggregation_failures()),
               "last_aggregation_duration_ms": 0.0, "last_aggregation_device_timed": 0.0,
               "last_round_duration_seconds": 0.0}
        for md in reversed(core.get_runtime_metadata_lineage(3)):
            blocks = md.get("model_aggregation_block_duration_ms") or []
            if blocks:
                out["last_aggregation_duration_ms"] = float(sum(blocks))
                out["last_aggregation_device_timed"] = 1.0 if md.get("device_timed") else 0.0
            if md.get("completed_at") and md.get("started_at"):
                out["last_round_duration_seconds"] = max(0.0, float(md["completed_at"]) - float(md["started_at"]))
            if blocks:
                break
        return out

    def start(self):
        if not self.enabled:
            return False
        try:
            from prometheus_client import CollectorRegistry, start_http_server
            from prometheus_client.core import CounterMetricFamily, GaugeMetricFamily
        except ImportError:
            MetisLogger.warning("METISFL_PROMETHEUS_PORT is set but prometheus_client is not installed")
            return False
        outer = self

        class _Collector:
            def collect(self):
                try:
                    snap = outer.snapshot()
                except Exception as e:   # a scrape must never disturb the controller
                    MetisLogger.warning("metrics scrape failed: %r", e)
                    return
                for key, value in snap.items():
                    name = "metisfl_" + key
                    if key.endswith("_total"):
                        m = CounterMetricFamily(name[:-len("_total")], key.replace("_", " "))
                        m.add_metric([], value)
                    else:
                        m = GaugeMetricFamily(name, key.replace("_", " "))
                        m.add_metric([], value)
                    yield m

        self.registry = CollectorRegistry()
        self.registry.register(_Collector())
        res = start_http_server(self._port, addr=self._addr, registry=self.registry)
        if isinstance(res, tuple):   # prometheus_client >= 0.17 returns (server, thread)
            self._httpd, self._thread = res
        MetisLogger.info("Prometheus metrics on %s:%d", self._addr, self._port)
        return True

    def stop(self):
        if self._httpd is not None:
            try:
                self._httpd.shutdown()
                self._httpd.server_close()
            except Exception:
                pass
            self._httpd = None
