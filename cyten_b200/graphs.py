"""CUDA-graph capture of a plan-cached operation chain.

Once the plans of a chain of hot-path operations are cached (same tensor structure every call: each Lanczos
iteration, each sweep step), a call is nothing but kernel launches, memsets and arena allocations -- no host-side
bookkeeping result is needed and no host<->device synchronisation happens.  That sequence is captured ONCE into a
CUDA graph and replayed: the ~0.4 ms of Python + launch overhead per H_eff matvec (which bounds the chi = 1024
matvec and the per-rank steps of the 8-GPU run) shrinks to one graph launch.

The input tensor's arena is static: the caller writes new block contents INTO ``step.input.data.buf`` (e.g. the
upload from pinned host memory) and calls ``step()``; the result tensor is static too and is overwritten by the
next replay.
"""
from __future__ import annotations

import torch


class CapturedStep:
    """``fn(input) -> tensor`` captured into a CUDA graph after ``warmup`` eager calls (which build the plans)."""

    def __init__(self, fn, input_tensor, warmup: int = 3):
        be = input_tensor.backend
        if be.device.type != "cuda":
            raise RuntimeError("CapturedStep needs a CUDA backend")
        self.input = input_tensor
        self.backend = be
        input_tensor.data.buf            # a pending permutation must be materialised before the capture
        for _ in range(max(1, warmup)):
            y = fn(input_tensor)
            y.data.buf
        torch.cuda.synchronize(be.device)
        # compose caches re-laid-out copies of an (immutable) operand arena on the operand; the static input is NOT
        # immutable -- its re-layout must be a node of the graph, not a buffer left over from the warm-up calls
        input_tensor.data._relayout = None
        misses = be.plan_misses
        saved = (be.gemm_events, be.plan_mkn)
        be.gemm_events, be.plan_mkn = None, None        # timing events cannot be recorded inside a capture
        self.graph = torch.cuda.CUDAGraph()
        try:
            with torch.cuda.device(be.device):
                launches = be.launches
                with torch.cuda.graph(self.graph):
                    y = fn(input_tensor)
                    y.data.buf                            # the final (lazy) permute_legs belongs to the graph
                self.launches_per_replay = be.launches - launches
        finally:
            be.gemm_events, be.plan_mkn = saved
        if be.plan_misses != misses:
            raise RuntimeError("a plan was built during graph capture: run more warm-up calls first")
        self.output = y
        input_tensor.data._relayout = None

    def __call__(self):
        self.graph.replay()
        self.backend.launches += self.launches_per_replay
        return self.output


class CapturedPipeline:
    """``parts[g](input)`` (side effects into static arenas, e.g. ``OverlappedMatvec.part``) captured as ONE GRAPH PER PART
    in a shared memory pool, so the caller can issue work that cannot live in the graphs -- the NCCL all_reduce of part
    g's range -- between the replays.  Later parts may read tensors an earlier part produced (same pool, kept alive by
    the producer's references; the replays run in order on one stream)."""

    def __init__(self, parts, input_tensor, warmup: int = 2):
        be = input_tensor.backend
        if be.device.type != "cuda":
            raise RuntimeError("CapturedPipeline needs a CUDA backend")
        self.input, self.backend = input_tensor, be
        input_tensor.data.buf
        for _ in range(max(1, warmup)):
            for fn in parts:
                fn(input_tensor)
        torch.cuda.synchronize(be.device)
        input_tensor.data._relayout = None          # see CapturedStep: the static input's re-layout is a graph node
        misses = be.plan_misses
        saved = (be.gemm_events, be.plan_mkn)
        be.gemm_events, be.plan_mkn = None, None
        self.graphs, self.launches = [], []
        pool = None
        try:
            with torch.cuda.device(be.device):
                for fn in parts:
                    g = torch.cuda.CUDAGraph()
                    n0 = be.launches
                    with torch.cuda.graph(g, pool=pool):
                        fn(input_tensor)
                    pool = g.pool() if pool is None else pool
                    self.graphs.append(g)
                    self.launches.append(be.launches - n0)
        finally:
            be.gemm_events, be.plan_mkn = saved
        if be.plan_misses != misses:
            raise RuntimeError("a plan was built during graph capture: run more warm-up calls first")
        self.launches_per_replay = sum(self.launches)
        input_tensor.data._relayout = None

    def replay(self, g: int):
        self.graphs[g].replay()
        self.backend.launches += self.launches[g]
