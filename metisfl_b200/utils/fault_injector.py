"""Fault injection for protocol tests (the reference has none).

Rules come from the ``METISFL_FAULTS`` environment variable (JSON list) or are added in code:
``{"learner": "<host:port or index>", "round": 3, "action": "drop" | "delay" | "kill", "seconds": 2.0}``.
``LearnerServicer.RunTask`` consults the injector before starting a task:
  drop  - acknowledge the task but never run it (a silent learner; exercises the controller's deadline
          eviction, quirk Q9),
  delay - sleep before training (a straggler; exercises semi-synchronous sizing / staleness),
  kill  - terminate the learner process (exercises membership change handling).
"""
import json
import os
import threading
import time


class FaultInjector(object):
    _instance = None
    _lock = threading.Lock()

    def __init__(self, rules=None):
        self.rules = list(rules or [])
        self.fired = []

    @classmethod
    def instance(cls):
        with cls._lock:
            if cls._instance is None:
                rules = []
                raw = os.environ.get("METISFL_FAULTS")
                if raw:
                    try:
                        rules = json.loads(raw)
                    except ValueError:
                        rules = []
                cls._instance = cls(rules)
            return cls._instance

    @classmethod
    def reset(cls, rules=None):
        with cls._lock:
            cls._instance = cls(rules or [])
            return cls._instance

    def add(self, learner, round, action, seconds=0.0):
        self.rules.append({"learner": str(learner), "round": int(round), "action": action, "seconds": float(seconds)})

    def on_task(self, learner_id: str, global_iteration: int) -> str:
        """Returns "run", or "drop" if the task must be swallowed; sleeps / exits for delay / kill."""
        for rule in self.rules:
            if str(rule.get("learner")) not in (learner_id, learner_id.split(":")[-1], "*"):
                continue
            if int(rule.get("round", -1)) not in (-1, int(global_iteration)):
                continue
            action = rule.get("action", "drop")
            self.fired.append((learner_id, global_iteration, action))
            if action == "delay":
                time.sleep(float(rule.get("seconds", 1.0)))
                return "run"
            if action == "kill":
                os._exit(17)
            return "drop"
        return "run"
