"""Restatement of the parts of fvcore.transforms.transform the TTA path uses (TEST INFRASTRUCTURE ONLY).

fvcore is a third-party dependency that is absent from /root/reference and from this image (`fvcore>=0.1.2`,
reference setup.py:214; unpinned beyond that), so these classes are restated from fvcore's published definitions:
"parity unpinned" for this file.  The reference's own code on the path (ResizeTransform in
detectron2/data/transforms/transform.py, GeneralizedRCNNWithTTAAVG._get_augmented_boxes) is executed in place on
top of these base classes by oracle/ref_loader.tta_reference().

  Transform.apply_box     corners ([x0,y0],[x1,y0],[x0,y1],[x1,y1]) -> apply_coords -> (min, max) per axis
  HFlipTransform          coords[:, 0] = width - coords[:, 0]
  NoOpTransform           identity coords (apply_box still runs the corner min / max)
  TransformList           applies its members in order; inverse() = reversed list of inverses; + concatenates
"""
import numpy as np


class Transform:
    def _set_attributes(self, params=None):
        if params:
            for k, v in params.items():
                if k != "self" and not k.startswith("_"):
                    setattr(self, k, v)

    def apply_coords(self, coords):
        raise NotImplementedError

    def apply_box(self, box):
        idxs = np.array([(0, 1), (2, 1), (0, 3), (2, 3)]).flatten()
        coords = np.asarray(box).reshape(-1, 4)[:, idxs].reshape(-1, 2)
        coords = self.apply_coords(coords).reshape((-1, 4, 2))
        minxy = coords.min(axis=1)
        maxxy = coords.max(axis=1)
        return np.concatenate((minxy, maxxy), axis=1)

    def inverse(self):
        raise NotImplementedError

    def __add__(self, other):
        return TransformList([self]) + other

    @classmethod
    def register_type(cls, data_type, func=None):
        # fvcore: adds an apply_<data_type> method; the reference registers "rotated_box" (not on this path)
        setattr(cls, "apply_" + data_type, func)


class NoOpTransform(Transform):
    def apply_image(self, img):
        return img

    def apply_coords(self, coords):
        return coords

    def inverse(self):
        return self


class HFlipTransform(Transform):
    def __init__(self, width):
        super().__init__()
        self._set_attributes(locals())

    def apply_coords(self, coords):
        coords[:, 0] = self.width - coords[:, 0]
        return coords

    def inverse(self):
        return self


class CropTransform(Transform):  # imported by the reference's transform.py; not on the path
    pass


class TransformList(Transform):
    def __init__(self, transforms):
        super().__init__()
        tfms = []
        for t in transforms:
            assert isinstance(t, Transform), t
            tfms.extend(t.transforms if isinstance(t, TransformList) else [t])
        self.transforms = tfms

    def apply_box(self, x):
        for t in self.transforms:
            x = t.apply_box(x)
        return x

    def apply_coords(self, x):
        for t in self.transforms:
            x = t.apply_coords(x)
        return x

    def __add__(self, other):
        others = other.transforms if isinstance(other, TransformList) else [other]
        return TransformList(self.transforms + others)

    def __radd__(self, other):
        others = other.transforms if isinstance(other, TransformList) else [other]
        return TransformList(others + self.transforms)

    def __len__(self):
        return len(self.transforms)

    def inverse(self):
        return TransformList([t.inverse() for t in self.transforms[::-1]])
