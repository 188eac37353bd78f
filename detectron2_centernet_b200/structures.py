"""The I/O contract of the ctdet path: minimal `Boxes` and `Instances` containers.

They carry exactly what the hot path reads and writes (detectron2/structures/boxes.py:133,
instances.py:7 in the reference): `Instances(image_size)` with tensor-like fields of equal
length, attribute access, `len`, indexing, `to`, `get_fields`; `Boxes.tensor [N,4]` XYXY fp32 with
`clip`, `nonempty`, `scale`.  When the real detectron2 is importable the wrappers in this package
accept its classes as well (duck typing: `.gt_boxes.tensor`, `.gt_classes`, `.image_size`).
"""
import torch


class Boxes:
    def __init__(self, tensor):
        if not isinstance(tensor, torch.Tensor):
            tensor = torch.as_tensor(tensor, dtype=torch.float32)
        tensor = tensor.to(torch.float32)
        if tensor.numel() == 0:
            tensor = tensor.reshape(0, 4)
        assert tensor.dim() == 2 and tensor.size(-1) == 4, tensor.size()
        self.tensor = tensor

    def __len__(self):
        return self.tensor.shape[0]

    def __getitem__(self, item):
        if isinstance(item, int):
            return Boxes(self.tensor[item].view(1, -1))
        return Boxes(self.tensor[item])

    def to(self, *args, **kwargs):
        return Boxes(self.tensor.to(*args, **kwargs))

    @property
    def device(self):
        return self.tensor.device

    def clone(self):
        return Boxes(self.tensor.clone())

    def area(self):
        t = self.tensor
        return (t[:, 2] - t[:, 0]) * (t[:, 3] - t[:, 1])

    def clip(self, box_size):
        assert torch.isfinite(self.tensor).all(), "Box tensor contains infinite or NaN!"
        h, w = box_size
        self.tensor[:, 0::2].clamp_(min=0, max=w)
        self.tensor[:, 1::2].clamp_(min=0, max=h)

    def nonempty(self, threshold=0.0):
        t = self.tensor
        return ((t[:, 2] - t[:, 0]) > threshold) & ((t[:, 3] - t[:, 1]) > threshold)

    def scale(self, scale_x, scale_y):
        self.tensor[:, 0::2] *= scale_x
        self.tensor[:, 1::2] *= scale_y

    def __repr__(self):
        return "Boxes(" + str(self.tensor) + ")"


class Instances:
    def __init__(self, image_size, **kwargs):
        object.__setattr__(self, "_image_size", tuple(
            int(x.item()) if isinstance(x, torch.Tensor) else x for x in image_size))
        object.__setattr__(self, "_fields", {})
        for k, v in kwargs.items():
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
        fields = object.__getattribute__(self, "_fields")
        if name not in fields:
            raise AttributeError("Cannot find field '{}' in the given Instances!".format(name))
        return fields[name]

    def set(self, name, value):
        n = len(value)
        if self._fields:
            assert len(self) == n, \
                "Adding a field of length {} to a Instances of length {}".format(n, len(self))
        self._fields[name] = value

    def has(self, name):
        return name in self._fields

    def get(self, name):
        return self._fields[name]

    def get_fields(self):
        return self._fields

    def remove(self, name):
        del self._fields[name]

    def to(self, *args, **kwargs):
        out = Instances(self._image_size)
        for k, v in self._fields.items():
            out.set(k, v.to(*args, **kwargs) if hasattr(v, "to") else v)
        return out

    def __getitem__(self, item):
        if isinstance(item, int):
            if item >= len(self) or item < -len(self):
                raise IndexError("Instances index out of range!")
            item = slice(item, None, len(self))
        out = Instances(self._image_size)
        for k, v in self._fields.items():
            out.set(k, v[item])
        return out

    def __len__(self):
        for v in self._fields.values():
            return len(v)
        raise NotImplementedError("Empty Instances does not support __len__!")

    def __repr__(self):
        return "Instances(num_instances={}, image_size={}, fields=[{}])".format(
            len(self) if self._fields else 0, self._image_size,
            ", ".join("{}: {}".format(k, v) for k, v in self._fields.items()))


class ImageList:
    """The slice of detectron2/structures/image_list.py:9 the ctdet path uses: the padded batch
    tensor [N,C,H,W] plus the original (h, w) of every image."""

    def __init__(self, tensor, image_sizes):
        self.tensor = tensor
        self.image_sizes = [tuple(int(v) for v in s) for s in image_sizes]

    def __len__(self):
        return len(self.image_sizes)

    def __getitem__(self, idx):
        h, w = self.image_sizes[idx]
        return self.tensor[idx, ..., :h, :w]          # image_list.py:43-54

    def to(self, *args, **kwargs):
        return ImageList(self.tensor.to(*args, **kwargs), self.image_sizes)

    @property
    def device(self):
        return self.tensor.device
