"""Client object with the reference's exact API and return conventions (C7).

``IMDBClient(fl.client.NumPyClient)`` — src/Servercase/server_IID_IMDB.py:155-179:

* ``get_parameters(config)`` -> every state_dict entry, in key order, as fp32 ndarrays (:161-162)
* ``set_parameters(list)``   -> zip keys <-> arrays, strict load (:164-167)
* ``fit(params, config)``    -> ``(params, len(trainloader), {})`` — the weight is the number of
  **batches**, not samples (:169-174)
* ``evaluate(params, config)`` -> ``(loss, len(testloader), {"accuracy", "loss"})`` (:176-179)
* serverless extras ``train_model()`` / ``evaluate_model()``
  (src/Serverlesscase/serverless_IID_IMDB.py:156-187)

The ndarray API is kept for parity; the framework's own round loops use ``get_flat`` /
``set_flat`` so weights never leave the device (the reference D2H-copies 433 MB per client
per round, serverless_IID_IMDB.py:263).
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np
import torch


class FedClient:
    def __init__(self, runner, trainloader, testloader, cid: int = 0, local_epochs: int = 1):
        self.net = runner
        self.trainloader = trainloader
        self.testloader = testloader
        self.cid = cid
        self.local_epochs = local_epochs

    # ---- Flower NumPyClient surface ---------------------------------------------------
    def get_parameters(self, config=None) -> List[np.ndarray]:
        return [v.detach().cpu().numpy().copy() for v in self.net.state_dict().values()]

    def set_parameters(self, parameters: List[np.ndarray]) -> None:
        keys = list(self.net.state_dict().keys())
        if len(keys) != len(parameters):
            raise ValueError(f"expected {len(keys)} tensors, got {len(parameters)}")
        self.net.load_state_dict({k: torch.as_tensor(np.asarray(v)) for k, v in zip(keys, parameters)})
        self.net.sync_shadow()

    def fit(self, parameters, config=None) -> Tuple[List[np.ndarray], int, Dict]:
        self.set_parameters(parameters)
        self.net.train_epochs(self.trainloader, self.local_epochs)
        return self.get_parameters({}), len(self.trainloader), {}

    def evaluate(self, parameters, config=None) -> Tuple[float, int, Dict[str, float]]:
        self.set_parameters(parameters)
        loss, acc = self.net.evaluate(self.testloader)
        return float(loss), len(self.testloader), {"accuracy": float(acc), "loss": float(loss)}

    # ---- serverless extras --------------------------------------------------------------
    def train_model(self) -> Dict[str, float]:
        return self.net.train_epochs(self.trainloader, self.local_epochs)

    def evaluate_model(self) -> Tuple[float, float]:
        return self.net.evaluate(self.testloader)

    # ---- device-resident fast path --------------------------------------------------------
    def get_flat(self) -> torch.Tensor:
        return self.net.flat

    def set_flat(self, flat: torch.Tensor) -> None:
        self.net.load_flat(flat)
        self.net.sync_shadow()

    def fit_flat(self, flat: torch.Tensor) -> Tuple[torch.Tensor, int, Dict]:
        self.set_flat(flat)
        stats = self.net.train_epochs(self.trainloader, self.local_epochs)
        return self.net.flat, len(self.trainloader), stats

    def evaluate_flat(self, flat: torch.Tensor) -> Tuple[float, int, Dict[str, float]]:
        self.set_flat(flat)
        loss, acc = self.net.evaluate(self.testloader)
        return float(loss), len(self.testloader), {"accuracy": float(acc), "loss": float(loss)}


# the reference's class name, for scripts that switch over
IMDBClient = FedClient
