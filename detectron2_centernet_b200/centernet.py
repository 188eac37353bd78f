"""Host-side mirror of the CenterNet meta-architecture's hot-path methods
(detectron2/modeling/meta_arch/centernet.py): `preprocess targets` (:186-189), `losses` (:191-212),
`inference` (:214-234) and `inference_single_image` (:236-266), re-expressed over whole batches and
backed by the sm_100a kernels.  The backbone / heads stay ordinary torch modules; this class takes
their outputs `{'hm','wh','reg'}` (names pinned by export/meta_modeling.py:200)."""
import numpy as np
import torch

from . import decode as D
from . import dist as ctdist
from . import losses as L
from . import preprocess as P
from . import target as T
from .config import get_cfg
from .structures import Boxes, Instances


class CenterNetHotPath:
    def __init__(self, cfg=None, num_classes=None, sync_num_pos=False):
        cfg = cfg or get_cfg()
        c = cfg["MODEL"]["CENTERNET"]
        # the reference overrides TASK.HM with len(meta.thing_classes) (:59-63)
        self.num_classes = int(num_classes if num_classes is not None else c["NUM_CLASSES"])
        self.down_ratio = int(c["DOWN_RATIO"])
        self.hm_weight, self.wh_weight, self.off_weight = float(c["HM_WEIGHT"]), float(c["WH_WEIGHT"]), float(c["OFF_WEIGHT"])
        self.focal_loss_alpha = list(c["FOCAL_LOSS_ALPHA"])
        self.score_threshold = float(c["SCORE_THRESH_TEST"])
        self.topk_candidates = int(c["TOPK_CANDIDATES_TEST"])
        self.max_detections_per_image = int(cfg["TEST"]["DETECTIONS_PER_IMAGE"])
        self.sync_num_pos = bool(sync_num_pos)
        self.pixel_mean = list(cfg["MODEL"]["PIXEL_MEAN"])                      # :64-65
        self.pixel_std = list(cfg["MODEL"]["PIXEL_STD"])
        self.size_divisibility = int(c["SIZE_DIVISIBILITY"])                  # :72 / :111

    # ---- input ---------------------------------------------------------------------------------
    def preprocess_image(self, batched_inputs, device=None):
        """:173-184: `x["image"] / 255`, normalise, pad and batch -> ImageList.  One launch for the batch."""
        images = [x["image"] if device is None else x["image"].to(device, non_blocking=True) for x in batched_inputs]
        return P.preprocess_image(images, self.pixel_mean, self.pixel_std, self.size_divisibility)

    # ---- training ------------------------------------------------------------------------------
    def targets(self, instances, input_shape):
        """`[gen_heatmap(x, output_shape, meta) for x in instances]` + the torch.cat of :197-201 in
        one launch.  input_shape = (H, W) of the padded batch."""
        out_shape = (int(input_shape[0]) // self.down_ratio, int(input_shape[1]) // self.down_ratio)
        return T.gen_heatmap_batch(instances, out_shape, self.num_classes)

    def losses(self, outputs, targets):
        """:191-212 -> {"hm_loss", "wh_loss", "off_loss"} (already weighted)."""
        norm = ctdist.sync_normalisers(targets) if self.sync_num_pos else None
        return L.ctdet_losses(outputs, targets, alpha=self.focal_loss_alpha,
                              weights=(self.hm_weight, self.wh_weight, self.off_weight), norm=norm)

    def losses_from_instances(self, outputs, instances):
        """`preprocess targets` (:188) + `losses` (:191-212) in ONE pass: the Gaussian heatmap target is
        generated inside the focal-loss sweep and never written to HBM.  Returns (loss dict, slot targets)."""
        if self.sync_num_pos:
            # global normalisers need the slot targets of every rank before the sweep: two-pass path
            hm = outputs["hm"]
            targets = T.gen_heatmap_batch(instances, (hm.shape[2], hm.shape[3]), self.num_classes, hm.device)
            return self.losses(outputs, targets), {k: v for k, v in targets.items() if k != "hm"}
        boxes, classes, offsets = T.pack_instances(instances, outputs["hm"].device)
        return L.ctdet_losses_from_objects(outputs, boxes, classes, offsets, alpha=self.focal_loss_alpha,
                                           weights=(self.hm_weight, self.wh_weight, self.off_weight))

    # ---- inference -----------------------------------------------------------------------------
    def decode(self, outputs):
        """Whole batch: sigmoid+clamp (:161) folded in (ranking on logits), top-K, boxes."""
        return D.ctdet_decode_batch(outputs["hm"], outputs["wh"], outputs["reg"], down_ratio=self.down_ratio,
                                    K=self.topk_candidates, logits=True)

    def inference(self, outputs, image_sizes, output_sizes=None):
        """:214-234 (+ detector_postprocess, modeling/postprocessing.py:10-74, when output_sizes is
        given) -> list of Instances.  One decode launch and one post-process launch for the batch;
        a single D2H copy of the per-image counts instead of a host sync per image (:258-261)."""
        boxes, scores, classes, _ = self.decode(outputs)
        B = scores.shape[0]
        out_sizes = image_sizes if output_sizes is None else output_sizes
        # without output_sizes this is the reference's `inference` alone: threshold + slice, boxes as decoded
        # (out_h = out_w = 0 tells the kernel to skip detector_postprocess' scale / clip / nonempty)
        sizes = torch.tensor([[int(i[0]), int(i[1]), int(o[0]), int(o[1])] if output_sizes is not None
                              else [int(i[0]), int(i[1]), 0, 0] for i, o in zip(image_sizes, out_sizes)],
                             dtype=torch.int32).to(scores.device, non_blocking=True)
        n_keep = D.postprocess_batch(boxes, scores, classes, sizes, self.max_detections_per_image,
                                     self.score_threshold).cpu().numpy()
        results = []
        for b in range(B):
            n = int(n_keep[b])
            r = Instances(tuple(int(v) for v in out_sizes[b]))
            r.pred_boxes = Boxes(boxes[b, :n])
            r.scores = scores[b, :n]
            r.pred_classes = classes[b, :n]
            results.append(r)
        return results

    def inference_single_image(self, output, image_size):
        """:236-266 for a batch-of-one output dict."""
        return self.inference(output, [image_size])[0]
