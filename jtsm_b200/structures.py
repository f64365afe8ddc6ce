"""Minimal `Boxes` / `Instances` containers with the attribute surface the ROI heads use
(`proposal_boxes.tensor`, `len()`, `has()`, `image_size`, field access by attribute, `Boxes.cat`).

The head modules of this package are duck-typed: inside a detectron2 installation they take detectron2's own
`Instances` / `Boxes` (detectron2/structures/instances.py, boxes.py:130-310) unchanged; these stand-ins exist because
detectron2 cannot be imported where this repo is built and tested, and for bench.py / the tests."""
import torch


class Boxes:
    def __init__(self, tensor):
        if not torch.is_tensor(tensor):
            tensor = torch.as_tensor(tensor, dtype=torch.float32)
        if tensor.numel() == 0:
            tensor = tensor.reshape(-1, 4).to(torch.float32)
        assert tensor.dim() == 2 and tensor.size(-1) == 4, tensor.size()
        self.tensor = tensor

    def __len__(self):
        return self.tensor.shape[0]

    def to(self, *a, **k):
        return Boxes(self.tensor.to(*a, **k))

    @property
    def device(self):
        return self.tensor.device

    def __getitem__(self, item):
        t = self.tensor[item]
        return Boxes(t.view(1, -1) if t.dim() == 1 else t)

    @classmethod
    def cat(cls, boxes_list):
        if len(boxes_list) == 0:
            return cls(torch.empty(0, 4))
        return cls(torch.cat([b.tensor for b in boxes_list], dim=0))


class Instances:
    def __init__(self, image_size, **fields):
        object.__setattr__(self, "_image_size", tuple(image_size))
        object.__setattr__(self, "_fields", {})
        for k, v in fields.items():
            self.set(k, v)

    @property
    def image_size(self):
        return self._image_size

    def __setattr__(self, name, val):
        if name.startswith("_"):
            object.__setattr__(self, name, val)
        else:
            self.set(name, val)

    def __getattr__(self, name):
        if name == "_fields" or name not in self._fields:
            raise AttributeError("Cannot find field '%s' in the given Instances!" % name)
        return self._fields[name]

    def set(self, name, value):
        n = len(value)
        if len(self._fields):
            assert len(self) == n, "Adding a field of length %d to Instances of length %d" % (n, len(self))
        self._fields[name] = value

    def has(self, name):
        return name in self._fields

    def get(self, name):
        return self._fields[name]

    def get_fields(self):
        return self._fields

    def __len__(self):
        for v in self._fields.values():
            return len(v)
        raise NotImplementedError("Empty Instances does not support __len__!")
