"""Shared helpers for the tests: golden-fixture loading."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


class Fixture:
    def __init__(self, name):
        self.z = np.load(os.path.join(GOLDEN, name + ".npz"))

    def __getitem__(self, k):
        return self.z[k]

    def t(self, k):
        return torch.from_numpy(self.z[k])

    def keys(self, prefix=""):
        return [k for k in self.z.files if k.startswith(prefix)]

    def state_dict(self):
        return {k[3:]: torch.from_numpy(self.z[k]) for k in self.z.files if k.startswith("sd/")}

    def stats(self, dtype=torch.float32):
        out = {}
        for k in self.z.files:
            if k.startswith("stats/"):
                _, a, b = k.split("/")
                out.setdefault(a, {})[b] = torch.from_numpy(self.z[k]).to(dtype)
        return out

    def grads(self):
        return {k[len("grad_f64/"):]: torch.from_numpy(self.z[k]) for k in self.keys("grad_f64/")}
