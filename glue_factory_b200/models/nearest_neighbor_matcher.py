"""Nearest-neighbour matcher for normalised descriptors on the B200 kernels.

Drop-in for `gluefactory.models.matchers.nearest_neighbor_matcher.NearestNeighborMatcher` (reference:
gluefactory/models/matchers/nearest_neighbor_matcher.py:39-72, studied for behaviour only): same
`default_conf`, same `required_data_keys`, same output dictionary.

  similarity      fp32 descriptors are split into bf16 (hi, lo) pairs and multiplied on the tcgen05 path as
                  hi.hi + hi.lo + lo.hi with fp32 accumulation (K = 3 D), i.e. fp32-grade dot products
                  (the dropped lo.lo term is ~2^-16 relative) -- the ratio and distance tests compare
                  2 (1 - sim), so a plain bf16 product would move matches.
  matches         one statistics pass over the similarity gives per row and per column the soft-max
                  normalisers, the nearest neighbour (lowest index on ties) and the runner-up; find_nn in both
                  directions, the mutual check and the 0/1 scores are a few bytes per keypoint after that.
  log_assignment  a second pass writes 2 sim - LSE_row - LSE_col; the dustbin row / column are zero.

There is no PyTorch fallback: CPU tensors raise.
"""
import torch

from gluefactory.models.base_model import BaseModel

from .. import ops


class NearestNeighborMatcher(BaseModel):
    default_conf = {
        "ratio_thresh": None,
        "distance_thresh": None,
        "mutual_check": True,
        "loss": None,
    }
    required_data_keys = ["descriptors0", "descriptors1"]

    def _init(self, conf):
        if conf.loss == "N_pair":  # keeps the reference's state-dict schema (:47-50)
            self.register_parameter("temperature", torch.nn.Parameter(torch.tensor(1.0)))

    @staticmethod
    def similarity(desc0, desc1):
        """einsum('bnd,bmd->bnm') (:52) for fp32 descriptors [B,M,D], [B,N,D]; D a multiple of 64."""
        if desc0.dim() != 3 or desc1.dim() != 3 or desc0.shape[0] != desc1.shape[0] or desc0.shape[2] != desc1.shape[2]:
            raise ValueError(f"descriptors must be [B,M,D] and [B,N,D], got {tuple(desc0.shape)} and {tuple(desc1.shape)}")
        B, M, D = desc0.shape
        N = desc1.shape[1]
        if D % 64 != 0:
            raise NotImplementedError(f"the B200 similarity kernel needs a descriptor dimension that is a multiple of 64 (D={D})")
        if M < 1 or N < 1:
            raise ValueError(f"nearest-neighbour matching needs at least one keypoint per image (M={M}, N={N})")
        dev = desc0.device
        d0 = desc0.detach().float().contiguous()
        d1 = desc1.detach().float().contiguous()
        T = B * (M + N)
        split = torch.empty(T, 3 * D, dtype=torch.bfloat16, device=dev)
        ops.split_hilo(d0, B * M, D, B * M, split)                # A side: [hi | hi | lo]
        ops.split_hilo(d1, B * N, D, 0, split[B * M:])            # B side: [hi | lo | hi]
        sim = torch.empty(B, M, N, dtype=torch.float32, device=dev)
        ops.gemm(split, split, M=M, N=N, K=3 * D, lda=3 * D, ldb=3 * D, a_rows=T, b_row0=B * M, b_rows=B * N, batch=B,
                 a_batch_rows=M, b_batch_rows=N, block_n=256 if N > 128 else 128, epilogue=ops.EPI_F32, alpha=1.0,
                 out32=sim, ld32=N, out32_batch_stride=M * N)
        return sim

    @ops.on_device("descriptors0")
    def _forward(self, data):
        c = self.conf
        if c.ratio_thresh and min(data["descriptors0"].shape[1], data["descriptors1"].shape[1]) < 2:
            raise RuntimeError("the ratio test needs at least two keypoints per image (topk(2) in the reference)")
        sim = self.similarity(data["descriptors0"], data["descriptors1"])
        la, m0, m1, ms0, ms1 = ops.nn_match(sim, c.ratio_thresh, c.distance_thresh, c.mutual_check)
        return {
            "matches0": m0,
            "matches1": m1,
            "matching_scores0": ms0,
            "matching_scores1": ms1,
            "similarity": sim,
            "log_assignment": la,
        }

    def loss(self, pred, data):
        raise NotImplementedError("training (the N_pair loss) is outside the forward-only scope of the B200 path")

    def metrics(self, pred, data):
        raise NotImplementedError
