"""Cluster description in the TF_CONFIG style — parity with tensorflow_impl/rsrcs/network.py:36-90:

    {"cluster": {"ps": ["host:port", ...], "worker": ["host:port", ...]},
     "task": {"type": "ps"|"worker", "index": 0, "strategy_gradient": "Krum", "strategy_model": "Median", "attack": "None"}}
"""

from __future__ import annotations

import json

__all__ = ["Network"]


class Network:
    def __init__(self, tf_location):
        """`tf_location`: path of the JSON file, a JSON string, or an already-parsed dict."""
        if isinstance(tf_location, dict):
            config = tf_location
        else:
            text = str(tf_location)
            config = json.loads(text) if text.lstrip().startswith("{") else json.load(open(text))
        self.config = config
        self.cluster = config["cluster"]
        self.task = config["task"]

    def get_task_type(self):
        return self.task["type"]

    def get_task_index(self):
        return int(self.task["index"])

    def get_model_strategy(self):
        return self.task.get("strategy_model", "Average")

    def get_gradient_strategy(self):
        return self.task.get("strategy_gradient", "Average")

    def get_my_attack(self):
        return self.task.get("attack", "None")

    def get_all_ps(self):
        return list(self.cluster.get("ps", []))

    def get_all_workers(self):
        return list(self.cluster.get("worker", []))

    def get_all_other_worker(self):
        mine = self.get_my_node() if self.get_task_type() == "worker" else None
        return [w for w in self.get_all_workers() if w != mine]

    def get_my_node(self):
        return self.cluster[self.get_task_type()][self.get_task_index()]

    def get_my_port(self):
        return int(self.get_my_node().rsplit(":", 1)[1])
