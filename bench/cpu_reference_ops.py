#!/usr/bin/env python
"""Timed CPU baseline of the reference's own model-exchange arithmetic (no GPU needed):

* ``FedAvg.aggregate`` (p2pfl/learning/aggregators/fedavg.py:43-60) over 8 contributors,
* ``LightningLearner.encode_parameters`` / ``decode_parameters`` (lightning_learner.py:113-138),

run from the UNMODIFIED package in baseline/_ref on state_dicts of the three model sizes of
BASELINE.json.  These are the host-side operations the fused exchange kernels replace; the numbers sit
next to the kernels' in BASELINE.md.  Prints one JSON line per model."""
import json, os, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "baseline", "shim"), os.path.join(ROOT, "baseline", "_ref")]
import torch  # noqa: E402
from p2pfl.learning.aggregators.fedavg import FedAvg  # noqa: E402
from p2pfl.learning.pytorch.lightning_learner import LightningLearner  # noqa: E402
from p2pfl_b200.models import MLP, ResNet18, gpt2_small  # noqa: E402


def best(fn, n):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3, sum(ts) / len(ts) * 1e3


for name, make, reps in (("mlp", lambda: MLP(seed=0), 10), ("resnet18", lambda: ResNet18(seed=0), 5), ("gpt2", lambda: gpt2_small(seed=0), 2)):
    model = make()
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    n_float = sum(v.numel() for v in sd.values() if v.is_floating_point())
    contributors = {f"n{i}": ({k: (v + i if v.is_floating_point() else v) for k, v in sd.items()}, 100 + i) for i in range(8)}
    agg = FedAvg(node_name="cpu-bench")
    fa = best(lambda: agg.aggregate(contributors), reps)
    learner = LightningLearner.__new__(LightningLearner)  # only the two codec methods are used
    learner.get_parameters = lambda: sd  # type: ignore[method-assign]
    blob = learner.encode_parameters(sd)
    enc = best(lambda: learner.encode_parameters(sd), reps)
    dec = best(lambda: learner.decode_parameters(blob), reps)
    print(json.dumps({
        "bench": "reference_cpu_ops", "model": name, "float_params": n_float, "payload_MB": len(blob) / 1e6,
        "cpu_threads": torch.get_num_threads(), "fedavg8_ms_best": fa[0], "fedavg8_ms_mean": fa[1],
        "encode_ms_best": enc[0], "decode_ms_best": dec[0],
        "one_exchange_8_nodes_ms": 7 * (enc[0] + dec[0]) / 7 + fa[0],
        "note": "per node and round the reference encodes once per peer (7x, Q13), decodes 7 payloads and aggregates 8 models; "
                "one_exchange = 1 encode + 1 decode + 1 aggregate, a lower bound",
    }), flush=True)
