"""Split one dataset into per-learner shards (reference ``examples/utils/data_partitioning.py:8-124``:
IID, k-classes-per-partition non-IID; the Dirichlet split it only declares is implemented here)."""
import json

import numpy as np


class DataPartitioning(object):

    def __init__(self, x_train, y_train, partitions_num, seed=1990):
        self.x_train, self.y_train = np.asarray(x_train), np.asarray(y_train).reshape(-1)
        self.partitions_num = int(partitions_num)
        self.rng = np.random.default_rng(seed)
        self.partition_policy = None
        self.partitions = None

    def _take(self, index_lists):
        self.partitions = [(self.x_train[idx], self.y_train[idx]) for idx in index_lists]
        return [p[0] for p in self.partitions], [p[1] for p in self.partitions]

    def iid_partition(self):
        self.partition_policy = "IID"
        perm = self.rng.permutation(len(self.y_train))
        return self._take(np.array_split(perm, self.partitions_num))

    def non_iid_partition(self, classes_per_partition=2):
        """Sort by label, cut into ``partitions * classes_per_partition`` shards, deal shards round-robin
        so that every learner sees about ``classes_per_partition`` classes."""
        self.partition_policy = "Non-IID-{}".format(classes_per_partition)
        order = np.argsort(self.y_train, kind="stable")
        shards = np.array_split(order, self.partitions_num * classes_per_partition)
        deal = self.rng.permutation(len(shards))
        groups = [np.concatenate([shards[s] for s in deal[i::self.partitions_num]])
                  for i in range(self.partitions_num)]
        return self._take(groups)

    def dirichlet_based_partition(self, a):
        self.partition_policy = "Dirichlet-{}".format(a)
        classes = np.unique(self.y_train)
        groups = [[] for _ in range(self.partitions_num)]
        for c in classes:
            idx = self.rng.permutation(np.where(self.y_train == c)[0])
            cuts = (np.cumsum(self.rng.dirichlet([a] * self.partitions_num)) * len(idx)).astype(int)[:-1]
            for g, part in zip(groups, np.split(idx, cuts)):
                g.extend(part.tolist())
        return self._take([np.array(g, dtype=np.int64) for g in groups])

    def to_json_representation(self):
        out = {"partition_policy": self.partition_policy, "partitions": []}
        for i, (_, y) in enumerate(self.partitions or []):
            cls, cnt = np.unique(y, return_counts=True)
            out["partitions"].append({"partition_id": i, "size": int(len(y)),
                                      "classes": {int(c): int(n) for c, n in zip(cls, cnt)}})
        return json.dumps(out)
