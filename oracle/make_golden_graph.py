"""Generates tests/golden/graph_utils.npz by running the reference's OWN index helpers
(hislam2/geom/graph_utils.py, hislam2/factor_graph.py FactorGraph edge methods) unmodified on CPU.
Build-container only (needs /root/reference); modules the file imports but the edge methods never
touch (lietorch, networkx, matplotlib, scipy, sklearn, tqdm, util.utils) are stubbed."""
import functools
import importlib
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("DBA_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "graph_utils.npz")


class _Stub(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Stub(self.__name__ + "." + name)

    def __call__(self, *a, **k):
        return None


def load_reference():
    for name in ["lietorch", "networkx", "matplotlib", "matplotlib.pyplot", "matplotlib.collections", "scipy",
                 "scipy.spatial", "scipy.spatial.transform", "sklearn", "sklearn.decomposition", "tqdm", "util",
                 "util.utils"]:
        try:
            importlib.import_module(name)
        except Exception:
            sys.modules[name] = _Stub(name)
    sys.path.insert(0, os.path.join(REF, "hislam2"))
    sys.path.insert(0, os.path.join(REF, "hislam2", "geom"))
    import graph_utils
    graph_utils.meshgrid = functools.partial(graph_utils.meshgrid, device="cpu")   # default device is 'cuda'
    import factor_graph
    return graph_utils, factor_graph


def main():
    gu, fg = load_reference()
    out = {}
    for n, r in [(8, 2), (12, 3), (5, 7), (1, 1)]:
        ii, jj = gu.neighbourhood_graph(n, r)
        out[f"nbr_{n}_{r}_ii"], out[f"nbr_{n}_{r}_jj"] = ii.numpy(), jj.numpy()
    ii, jj = gu.meshgrid(3, 5)
    out["mesh_3_5_ii"], out["mesh_3_5_jj"] = ii.numpy(), jj.numpy()
    graph = {4: [1, 2], 0: [3], 7: [], 2: [0, 4, 7]}
    a, b, c = gu.graph_to_edge_list(graph)
    out["g2e_ii"], out["g2e_jj"], out["g2e_kk"] = a.numpy(), b.numpy(), c.numpy()
    out["kfi"] = gu.keyframe_indicies(graph).numpy()

    # FactorGraph edge methods: a scripted sequence of calls, state recorded after every step
    kf = types.SimpleNamespace(tstamp=torch.zeros(64), image=torch.zeros(64), pose=torch.zeros(64), intrinsic=torch.zeros(64),
                               get_lock=lambda: __import__("contextlib").nullcontext())
    G = fg.FactorGraph(kf, device="cpu", max_factors=48)
    gen = torch.Generator().manual_seed(7)
    step = 0

    def rec():
        nonlocal step
        out[f"fg_{step}_ii"], out[f"fg_{step}_jj"], out[f"fg_{step}_age"] = G.ii.numpy().copy(), G.jj.numpy().copy(), G.age.numpy().copy()
        step += 1

    script = []
    G.add_neighborhood_factors(0, 6, r=2); script.append(("nbr", 0, 6, 2)); rec()
    G.add_factors(torch.tensor([0, 5, 1, 5, 9]), torch.tensor([1, 0, 4, 0, 2])); script.append(("add", [0, 5, 1, 5, 9], [1, 0, 4, 0, 2], 0)); rec()
    G.age += torch.randint(0, 1000, G.age.shape, generator=gen); out["fg_age_bump0"] = G.age.numpy().copy(); script.append(("age", 0)); rec()
    G.add_neighborhood_factors(4, 12, r=3); script.append(("nbr", 4, 12, 3)); rec()
    G.age += torch.randint(0, 1000, G.age.shape, generator=gen) * 1000 + torch.arange(G.age.numel()); out["fg_age_bump1"] = G.age.numpy().copy(); script.append(("age", 1)); rec()
    ni, nj = [20, 21, 22, 20, 3, 21, 23, 24, 25, 26], [21, 20, 20, 22, 20, 3, 20, 20, 20, 20]
    G.add_factors(torch.tensor(ni), torch.tensor(nj), remove=True); script.append(("add", ni, nj, 1)); rec()
    G.rm_keyframe(5); script.append(("rmkf", 5)); rec()
    G.rm_factors(G.ii == 7); script.append(("rm_ii_eq", 7)); rec()
    G.rm_keyframe(0); script.append(("rmkf", 0)); rec()
    G.clear_edges(); script.append(("clear",)); rec()
    out["script"] = np.array(repr(script))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, "steps", step)


if __name__ == "__main__":
    main()
