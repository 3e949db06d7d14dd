"""SpeciesClassifier -- drop-in for the data path of the reference's src/species_classifier.py.

What is rebuilt here is what north_star names for Stage 2: the crop -> [N,3,S,S] gather
(`preprocess`, `crop_batch`: BGR->RGB, bilinear resize with the ATen tap rule, /255, ImageNet
mean/std; reference :298-352) as ONE batched CUDA kernel for all crops of a frame
(csrc/crop_resize.cu), and the result logic around the network (softmax / top-k / hierarchical
taxonomy label / geographic whitelist; reference :179-236, :354-474).  The classifier BODY is a
third-party timm model in the reference (:258-262), row N2 of SURVEY.md section 8(f).  For the family
the shipped configuration selects (`timm/eva02_large_patch14_clip_336...`, config/config.yaml:279-281;
any `eva02_*` name) `load_model` builds the network as a libtcd layer program (eva02_engine.py:
tcgen05 linears + hand-written attention / SwiGLU / softmax / top-k kernels) from the checkpoint's
state dict (`checkpoint_path`, or `weights=` in memory), and `classify_crops` feeds it patch rows
gathered straight from the frame.  Other timm names have no native body: a caller-supplied
`torch.nn.Module` (`model=` / `model_factory=`) mapping [N,3,S,S] -> logits is driven as the reference
drives its model, and without one `load_model` fails loudly (timm is not part of this image).
"""
import json
import logging
import time
from pathlib import Path
from typing import Any, Callable, Dict, List, Optional, Tuple

import numpy as np

logger = logging.getLogger(__name__)


class SpeciesClassifier:
    def __init__(self, model_name: str = "tf_efficientnet_b4_ns", checkpoint_path: Optional[str] = None,
                 device: str = "cuda", confidence_threshold: float = 0.3, taxonomy_file: Optional[str] = None,
                 input_size: int = 224, use_hierarchical: bool = True, allowed_species: Optional[List[str]] = None,
                 enable_geographic_filter: bool = False, model: Any = None,
                 model_factory: Optional[Callable[[int], Any]] = None, weights: Optional[Dict[str, Any]] = None,
                 max_batch: int = 64, batch_buckets: Tuple[int, ...] = (4, 16)):
        self.model_name = model_name
        self.checkpoint_path = checkpoint_path
        self.device = device
        self.confidence_threshold = confidence_threshold
        self.model = model
        self._factory = model_factory
        self._weights = weights          # in-memory state dict (timm key names) for the native EVA02 body
        self.max_batch = int(max_batch)  # crops per replay of the native body
        self.engine = None               # eva02_engine.Eva02Engine (max_batch slots) once load_model built it
        # A replay of a layer program computes every crop slot it was built for, so a frame with two detections should
        # not pay for 64: smaller programs (own weights, own CUDA graph) are built on first use for these sizes.
        self.batch_buckets = tuple(sorted(b for b in batch_buckets if 0 < b < self.max_batch))
        self._engines: Dict[int, Any] = {}
        self._sd = None
        self.taxonomy: Dict[Any, Any] = {}
        self.use_hierarchical = use_hierarchical
        self.enable_geographic_filter = enable_geographic_filter
        self.allowed_species = set(allowed_species) if allowed_species else set()
        self.input_size = input_size
        self.mean = [0.485, 0.456, 0.406]
        self.std = [0.229, 0.224, 0.225]
        self.hierarchy_thresholds = {"species": 0.5, "genus": 0.4, "family": 0.3, "order": 0.3, "class": 0.2}
        if taxonomy_file:
            self._load_taxonomy(taxonomy_file)

    # ------------------------------------------------------------------ taxonomy (reference :112-236)
    def _load_taxonomy(self, taxonomy_file: str, require_taxonomy: bool = False):
        try:
            path = Path(taxonomy_file)
            if not path.exists():
                raise FileNotFoundError("Taxonomy file not found: %s" % taxonomy_file)
            if path.suffix == ".json":
                with open(path) as f:
                    data = json.load(f)
                if not isinstance(data, dict):
                    raise ValueError("taxonomy JSON must be an object keyed by class id")
                self.taxonomy = data
            else:
                with open(path) as f:
                    self.taxonomy = {str(i): line.strip() for i, line in enumerate(f) if line.strip()}
        except Exception as e:
            if require_taxonomy:
                raise RuntimeError("Failed to load taxonomy from %s: %s" % (taxonomy_file, e))
            logger.warning("%s - continuing with generic labels", e)
            self.taxonomy = {}

    def get_hierarchical_label(self, class_id: int, confidence: float) -> Tuple[Optional[str], Optional[str]]:
        entry = self.taxonomy.get(str(class_id), {})
        if isinstance(entry, str):
            return entry, "species"
        if not self.use_hierarchical:
            if isinstance(entry, dict):
                return (entry.get("common_name") or entry.get("name") or "species_%d" % class_id), "species"
            return "species_%d" % class_id, "species"
        th = self.hierarchy_thresholds
        if confidence >= th["species"]:
            return (entry.get("common_name") or entry.get("name", "species_%d" % class_id)), "species"
        # one level per confidence band; an empty field in that band yields no label (reference :219-235)
        for level in ("genus", "family", "order", "class"):
            if confidence >= th[level]:
                val = entry.get(level, "")
                return (val, level) if val else (None, None)
        return None, None

    # ------------------------------------------------------------------ model
    _ENGINE_TOPK = 32   # the program's read-out width; classify(top_k <= 32) slices it

    def _load_native(self, num_classes: int):
        """EVA02 body as a libtcd layer program (replaces timm.create_model + load_state_dict, reference :258-268)."""
        from . import eva02_engine
        sd = self._weights
        if sd is None:
            if not self.checkpoint_path:
                raise RuntimeError("%s: pretrained weights come from the HuggingFace hub in the reference "
                                   "(pretrained=True); give checkpoint_path= or weights= here" % self.model_name)
            sd = self._read_checkpoint(self.checkpoint_path)
        sd = {k: (v.detach().cpu().float().numpy() if hasattr(v, "detach") else np.asarray(v, np.float32))
              for k, v in sd.items()}
        if sd["head.weight"].shape[0] != num_classes:
            raise ValueError("checkpoint head has %d classes, load_model asked for %d" % (sd["head.weight"].shape[0], num_classes))
        dev = str(self.device)
        idx = int(dev.split(":")[1]) if ":" in dev else 0
        self._sd, self._dev_idx = sd, idx
        self.engine = eva02_engine.Eva02Engine(sd, batch=self.max_batch, img_size=int(self.input_size),
                                               top_k=self._ENGINE_TOPK, device=idx)
        self._engines = {self.max_batch: self.engine}
        self.model = self.engine
        return True

    def _engine_for(self, n: int):
        """The smallest program that holds n crops (built on first use; the full-size one always exists)."""
        from . import eva02_engine
        size = next((b for b in self.batch_buckets if b >= n), self.max_batch)
        if size not in self._engines:
            self._engines[size] = eva02_engine.Eva02Engine(self._sd, batch=size, img_size=int(self.input_size),
                                                           top_k=self._ENGINE_TOPK, device=self._dev_idx)
        return self._engines[size]

    def close(self):
        for e in self._engines.values():
            e.close()
        self._engines, self.engine, self._sd = {}, None, None

    @staticmethod
    def _read_checkpoint(path: str) -> Dict[str, Any]:
        """State dict of a checkpoint file: a torch pickle as the reference loads it (`model_state_dict` or the dict
        itself, reference :265-267), or the `model.safetensors` file timm's hub models ship as."""
        if str(path).endswith(".safetensors"):
            from safetensors.numpy import load_file
            return dict(load_file(str(path)))
        import torch
        ck = torch.load(path, map_location="cpu")
        sd = ck["model_state_dict"] if "model_state_dict" in ck else ck
        return sd.get("state_dict", sd) if isinstance(sd, dict) else sd

    def load_model(self, num_classes: int = 1000, max_retries: int = 3) -> bool:
        from .eva02_engine import is_eva02
        if self.engine is not None:
            return True
        if self.model is None and self._factory is None and is_eva02(self.model_name):
            try:
                return self._load_native(num_classes)
            except Exception as e:
                logger.error("Failed to load species classifier: %s", e, exc_info=True)
                return False
        for attempt in range(max_retries):
            try:
                import torch
                if self.model is None:
                    if self._factory is not None:
                        self.model = self._factory(num_classes)
                    else:
                        import timm  # the reference's dependency; not part of this image
                        name = self.model_name
                        if name.startswith("timm/") or "/" in name:
                            name = "hf-hub:" + name
                        self.model = timm.create_model(name, pretrained=(self.checkpoint_path is None),
                                                       num_classes=num_classes)
                    if self.checkpoint_path:
                        ck = torch.load(self.checkpoint_path, map_location=self.device)
                        self.model.load_state_dict(ck["model_state_dict"] if "model_state_dict" in ck else ck)
                self.model = self.model.to(self.device).eval()
                return True
            except (RuntimeError, OSError, IOError, ConnectionError) as e:
                if attempt < max_retries - 1:
                    logger.warning("Model load failed (attempt %d/%d): %s", attempt + 1, max_retries, e)
                    time.sleep(2 ** attempt)
                else:
                    logger.error("Failed to load species classifier after %d attempts: %s", max_retries, e)
                    return False
            except Exception as e:
                logger.error("Failed to load species classifier: %s", e, exc_info=True)
                return False
        return False

    # ------------------------------------------------------------------ crop gather (the CUDA part)
    def _dev_frame(self, image):
        import torch
        t = image if isinstance(image, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(image))
        if not t.is_cuda:
            t = t.to(self.device if str(self.device).startswith("cuda") else "cuda")
        return t.contiguous()

    def crop_batch(self, frame, boxes):
        """All crops of one frame in one launch: boxes int [N,4] (x1,y1,x2,y2), already padded and
        clamped -> float32 CUDA tensor [N,3,S,S], normalised like `preprocess`."""
        import torch
        from . import engine
        t = self._dev_frame(frame)
        boxes = np.ascontiguousarray(boxes, np.int32).reshape(-1, 4)
        S = int(self.input_size)
        out = torch.empty((boxes.shape[0], 3, S, S), dtype=torch.float32, device=t.device)
        if boxes.shape[0]:
            engine.crop_resize(t.data_ptr(), int(t.shape[0]), int(t.shape[1]), t.stride(0), boxes, S, out.data_ptr(),
                               layout=0, swap_rb=1, scale=1.0 / 255.0, mean=self.mean, std=self.std)
        return out

    def preprocess(self, image):
        """(h,w,3) u8 BGR crop -> (1,3,S,S) float32 on the GPU (reference :298-352)."""
        h, w = int(image.shape[0]), int(image.shape[1])
        return self.crop_batch(image, [[0, 0, w, h]])

    # ------------------------------------------------------------------ results
    def _rows_to_results(self, probs_row, idx_row) -> List[Dict[str, Any]]:
        out = []
        floor = 0.1 if self.use_hierarchical else self.confidence_threshold
        for prob, idx in zip(probs_row, idx_row):
            prob, idx = float(prob), int(idx)
            if prob < floor:
                continue
            label, level = self.get_hierarchical_label(idx, prob)
            if label is None:
                continue
            if self.enable_geographic_filter and self.allowed_species and label not in self.allowed_species:
                continue
            out.append({"species": label, "confidence": prob, "class_id": idx, "taxonomic_level": level})
        return out

    def _engine_topk(self, eng, n, top_k):
        eng.run()
        eng.sync()
        p, i = eng.read_topk(n)
        k = min(top_k, p.shape[1])
        return p[:, :k], i[:, :k]

    def _topk(self, batch, top_k):
        import torch
        if self.engine is not None:
            # preprocessed [n,3,S,S] float32 CUDA tensor -> patch rows -> program replay, max_batch crops at a time
            batch = batch.contiguous().float()
            torch.cuda.current_stream(batch.device).synchronize()
            ps, is_ = [], []
            for s0 in range(0, batch.shape[0], self.max_batch):
                part = batch[s0:s0 + self.max_batch]
                eng = self._engine_for(part.shape[0])
                eng.patchify_device(part.data_ptr(), part.shape[0])
                p, i = self._engine_topk(eng, part.shape[0], top_k)
                ps.append(p)
                is_.append(i)
            return np.concatenate(ps, 0), np.concatenate(is_, 0)
        with torch.no_grad():
            probs = torch.softmax(self.model(batch), dim=1)
            p, i = torch.topk(probs, min(top_k, probs.shape[1]), dim=1)
        return p.cpu().numpy(), i.cpu().numpy()  # ONE sync for the whole batch (the reference .item()s per value)

    def classify(self, image, top_k: int = 5) -> List[Dict[str, Any]]:
        if self.model is None:
            logger.error("Model not loaded. Call load_model() first.")
            return []
        try:
            p, i = self._topk(self.preprocess(image), top_k)
            return self._rows_to_results(p[0], i[0])
        except Exception as e:
            logger.error("Classification failed: %s", e)
            return []

    def classify_crops(self, frame, boxes, top_k: int = 5) -> List[List[Dict[str, Any]]]:
        """Batched Stage 2: every box of `frame` gathered by one kernel, one forward, one D2H.
        Per-crop results equal `classify(frame[y1:y2, x1:x2], top_k)`."""
        if self.model is None:
            logger.error("Model not loaded. Call load_model() first.")
            return [[] for _ in boxes]
        if len(boxes) == 0:
            return []
        if self.engine is not None:
            # native body: the gather kernel writes the patch-embedding GEMM's operand directly (no NCHW tensor)
            import torch
            t = self._dev_frame(frame)
            torch.cuda.current_stream(t.device).synchronize()
            boxes = np.ascontiguousarray(boxes, np.int32).reshape(-1, 4)
            out = []
            for s0 in range(0, boxes.shape[0], self.max_batch):
                part = boxes[s0:s0 + self.max_batch]
                eng = self._engine_for(part.shape[0])
                eng.crop_patches(t.data_ptr(), int(t.shape[0]), int(t.shape[1]), t.stride(0), part, mean=self.mean, std=self.std)
                p, i = self._engine_topk(eng, part.shape[0], top_k)
                out.extend(self._rows_to_results(p[n], i[n]) for n in range(part.shape[0]))
            return out
        p, i = self._topk(self.crop_batch(frame, boxes), top_k)
        return [self._rows_to_results(p[n], i[n]) for n in range(len(boxes))]

    def classify_batch(self, images: List[np.ndarray], top_k: int = 5) -> List[List[Dict[str, Any]]]:
        """Reference :428-474: plain taxonomy lookup by int id, global threshold, no hierarchy."""
        if self.model is None:
            logger.error("Model not loaded. Call load_model() first.")
            return []
        try:
            import torch
            p, i = self._topk(torch.cat([self.preprocess(img) for img in images], 0), top_k)
            return [[{"species": self.taxonomy.get(int(c), "species_%d" % int(c)), "confidence": float(q),
                      "class_id": int(c)} for q, c in zip(p[n], i[n]) if float(q) >= self.confidence_threshold]
                    for n in range(len(images))]
        except Exception as e:
            logger.error("Batch classification failed: %s", e)
            return []
