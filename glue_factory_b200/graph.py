"""CUDA-graph replay of a fixed-shape forward pass.

A pair of views costs ~116 kernel launches through ctypes; at the benchmark's 32 pairs per step the GPU is the bound
(26 ms of kernels per step), but a serving loop with one or two pairs per call is launch-bound on the host.  Every entry
of the C ABI only enqueues work on the caller's stream (no allocation, no synchronisation, tensor maps are kernel
arguments), so a whole forward can be captured once and replayed with one launch.

No counterpart in the reference (its eager modules are re-launched op by op on every call).  Restrictions are those of
stream capture: fixed input shapes and dtypes, no data-dependent host decisions inside the pass (SuperPoint with
`max_num_keypoints` set and `detection_threshold < 0`, non-adaptive LightGlue / SuperGlue: the benchmarked
configurations; the adaptive path reads its final sizes on the host and is refused by the capture).
"""
import torch


def _map(x, fn):
    if torch.is_tensor(x):
        return fn(x)
    if isinstance(x, dict):
        return {k: _map(v, fn) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_map(v, fn) for v in x)
    return x


def _copy_into(dst, src, path="data"):
    if torch.is_tensor(dst):
        if not torch.is_tensor(src) or src.shape != dst.shape or src.dtype != dst.dtype:
            raise ValueError(f"{path}: the captured pass expects {tuple(dst.shape)} {dst.dtype}, got "
                             f"{tuple(src.shape) if torch.is_tensor(src) else type(src)} {getattr(src, 'dtype', '')}")
        dst.copy_(src, non_blocking=True)
    elif isinstance(dst, dict):
        for k, v in dst.items():
            if k not in src:
                raise KeyError(f"{path}[{k!r}] is missing")
            _copy_into(v, src[k], f"{path}[{k!r}]")
    elif isinstance(dst, (list, tuple)):
        for i, v in enumerate(dst):
            _copy_into(v, src[i], f"{path}[{i}]")


class GraphedForward:
    """`g = GraphedForward(model, example_data)`; `pred = g(data)` replays the captured pass on `data`.

    The returned tensors are the capture's own output buffers: the next call overwrites them (`clone=True` returns
    copies).  Inputs are copied into the capture's input buffers on the current stream; host tensors should be pinned.
    """

    def __init__(self, model, example, warmup=2, clone=False):
        dev = next(t for t in _flatten(example) if t.is_cuda).device
        self.model, self.clone, self.device = model, clone, dev
        self.static_in = _map(example, lambda t: t.to(dev).clone())
        with torch.cuda.device(dev):
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side), torch.no_grad():  # caches (weight packs, workspaces, sequence tables) fill here
                for _ in range(max(1, warmup)):
                    model(self.static_in)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph), torch.no_grad():
                self.static_out = model(self.static_in)

    def __call__(self, data):
        with torch.cuda.device(self.device):
            _copy_into(self.static_in, data)
            self.graph.replay()
        return _map(self.static_out, lambda t: t.clone()) if self.clone else self.static_out


def _flatten(x):
    if torch.is_tensor(x):
        yield x
    elif isinstance(x, dict):
        for v in x.values():
            yield from _flatten(v)
    elif isinstance(x, (list, tuple)):
        for v in x:
            yield from _flatten(v)
