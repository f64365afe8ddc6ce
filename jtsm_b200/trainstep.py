"""A data-parallel TRAINING step of the WSL ROI heads on the sm_100a path, for bench.py's `train_images_per_sec`.

What the reference runs per iteration after the backbone (projects/WSL/tools/train_net.py:74-119 ->
GeneralizedRCNNWSL.forward -> OICRROIHeads.forward, roi_heads_oicr.py:302-468), wrapped the way the reference wraps
its model (`DistributedDataParallel(model, device_ids=[local_rank], broadcast_buffers=False)`,
detectron2/engine/defaults.py:289-290):
    RoIPool (x objectness) -> fc6 -> fc7 (DiscriminativeAdaptionNeck, torch GEMMs; VGG16: 25088 -> 4096 -> 4096,
    ~120 M parameters) -> WSDDN MIL head + BCE -> Kr x (pseudo-gt, label assignment, weighted CE[, box regression])
    -> loss.backward() through the autograd Functions of roi_heads/functions.py (DDP all-reduces the gradient
    buckets over NCCL while the backward runs) -> SGD step.
The conv5 feature map is resident on the device with requires_grad (the backbone's output), so the pooling backward
runs too.  One process per GPU, each with its own seeded synthetic images."""
import torch
from torch import nn

from .box_regression import Box2BoxTransform
from .poolers import ROIPooler
from .roi_heads import DiscriminativeAdaptionNeck, OICROutputLayers, OICRROIHeadsB200, WSDDNOutputLayers
from .structures import Boxes, Instances


def build_heads(cfg, fc_dims=(4096, 4096)):
    """OICRROIHeadsB200 for a synth.CONFIGS entry (VGG16-DC5-like: one feature level, RoIPool P x P)."""
    k, kr = cfg["K"], cfg["Kr"]
    reg = [bool(cfg.get("reg"))] * max(kr, 1)
    shape = type("ShapeSpec", (), dict(channels=cfg["C"], height=cfg["P"], width=cfg["P"], stride=None))()
    pooler = ROIPooler(output_size=cfg["P"], scales=(cfg["scale"],), sampling_ratio=0, pooler_type="ROIPool")
    box_head = DiscriminativeAdaptionNeck(shape, fc_dims=tuple(fc_dims))
    b2b = Box2BoxTransform(weights=(10.0, 10.0, 5.0, 5.0))
    pred = WSDDNOutputLayers(box_head.output_shape, box2box_transform=b2b, num_classes=k, mean_loss=True)
    refs = [OICROutputLayers(box_head.output_shape, box2box_transform=b2b, num_classes=k, refine_k=i, refine_reg=reg,
                             box_reg_loss_type="smooth_l1_weighted" if cfg.get("reg") else "smooth_l1",
                             loss_weight={"loss_box_reg": 1.0}) for i in range(kr)]
    heads = OICRROIHeadsB200(num_classes=k, box_in_features=["conv5"], box_pooler=pooler, box_head=box_head,
                             box_predictor=pred, refine_K=kr, refine_reg=reg[:kr] if kr else [], box_refinery=refs)
    if not cfg.get("reg"):  # bbox_pred exists (checkpoint compatibility) but takes no part without WSL.REFINE_REG
        for r in refs:
            for p in r.bbox_pred.parameters():
                p.requires_grad_(False)
    return heads


class TrainStep:
    def __init__(self, cfg, device, world=1, fc_dims=(4096, 4096), lr=1e-3, tf32=False, fused_sgd=False):
        self.cfg, self.dev, self.world = cfg, device, world
        if tf32:  # the fc GEMMs are torch's; TF32 tensor cores keep fp32 storage / fp32 accumulate (process-wide switch)
            torch.backends.cuda.matmul.allow_tf32 = True
        torch.manual_seed(1234)  # identical initial replicas on every rank
        self.heads = build_heads(cfg, fc_dims).to(device)
        self.heads.train()
        self.model = self.heads
        if world > 1:
            # constructed on a side stream: what PyTorch asks for when the backward is later captured into a CUDA graph
            side = torch.cuda.Stream(device)
            side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(side):
                self.model = nn.parallel.DistributedDataParallel(self.heads, device_ids=[device.index],
                                                                 broadcast_buffers=False, gradient_as_bucket_view=True)
            torch.cuda.current_stream(device).wait_stream(side)
        self.params = [p for p in self.heads.parameters() if p.requires_grad]
        self.opt = torch.optim.SGD(self.params, lr=lr, momentum=0.9, weight_decay=5e-4, fused=bool(fused_sgd))
        self.graph = None

    def num_params(self):
        return sum(p.numel() for p in self.params)

    def make_batch(self, wl):
        """Device-resident inputs of one iteration from a synth.step_case workload."""
        cfg = self.cfg
        feat = wl["feat"].to(self.dev).requires_grad_()
        proposals, targets, off = [], [], 0
        for b, l in zip(wl["boxes"], wl["labels"]):
            n = b.shape[0]
            proposals.append(Instances((cfg["img_h"], cfg["img_w"]), proposal_boxes=Boxes(b.to(self.dev)),
                                       objectness_logits=wl["objectness"][off:off + n].to(self.dev)))
            targets.append(Instances((cfg["img_h"], cfg["img_w"]), gt_classes=l.to(self.dev)))
            off += n
        # image-level labels ready on the device (get_image_level_gt's torch.unique would synchronise every step)
        from . import ops
        from .roi_heads import get_image_level_gt

        gt_img, gt_int, oh = get_image_level_gt(targets, cfg["K"])
        lists = ops.image_class_lists([g.cpu() for g in gt_int], self.dev)
        return dict(feat=feat, proposals=proposals, targets=targets, image_level_gt=(gt_img, lists, oh))

    def step(self, batch, sync=True):
        """forward + backward (+ DDP all-reduce) + optimizer step -> total loss (device scalar)."""
        self.opt.zero_grad(set_to_none=True)
        batch["feat"].grad = None
        ctx = self.model.no_sync() if (self.world > 1 and not sync) else _null()
        with ctx:
            _, losses = self.model(None, {"conv5": batch["feat"]}, batch["proposals"], batch["targets"],
                                   image_level_gt=batch.get("image_level_gt"))
            loss = sum(losses.values())
            loss.backward()
        self.opt.step()
        self.last_losses = {k: v.detach() for k, v in losses.items()}  # detached: no autograd graph is kept alive
        return loss.detach()

    # ------------------------------------------------------------------ the whole step as ONE CUDA graph
    def capture(self, batch, warmup=11):
        """forward + backward (+ DDP all-reduce) + optimizer step captured into a CUDA graph over the STATIC tensors of
        `batch` (copy new inputs into them, then `replay()`).  Everything between the Linear layers goes through the C
        ABI on the capturing stream and nothing synchronises (the image-level labels come precomputed), so the step's
        ~105 launches cost one graph launch.  PyTorch's recipe for whole-network capture: eager warm-up iterations on a
        side stream (>= 11 with DDP), gradients set to None before the capture so that they live in the graph's pool."""
        dev = self.dev
        self.last_losses = None  # an autograd graph kept alive from an earlier stream would pin its AccumulateGrad nodes
        try:  # parameters that stepped eagerly on another stream keep their AccumulateGrad nodes: harmless here, noisy
            torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch(False)
        except AttributeError:
            pass
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup if self.world > 1 else 3):
                self.step(batch)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.opt.zero_grad(set_to_none=True)
        batch["feat"].grad = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            _, losses = self.model(None, {"conv5": batch["feat"]}, batch["proposals"], batch["targets"],
                                   image_level_gt=batch.get("image_level_gt"))
            loss = sum(losses.values())
            loss.backward()
            self.opt.step()
            self.graph_loss = loss.detach()
            del loss, losses
        self.graph_batch = batch
        return self.graph

    def replay(self):
        self.graph.replay()
        return self.graph_loss


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
