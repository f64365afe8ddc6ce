"""Shapes of the test-time augmentations without touching the GPU library (synth.py uses it on the host):
DatasetMapperTTAAVG.__call__ (test_time_augmentation_avg.py:147-196) resizes the image's short edge to every
MIN_SIZE (ResizeShortestEdge.get_transform, augmentation_impl.py:152-172) and adds a flipped copy of each."""


def resize_shortest_edge_shape(h, w, size, max_size):
    scale = size * 1.0 / min(h, w)
    if h < w:
        newh, neww = size, scale * w
    else:
        newh, neww = scale * h, size
    if max(newh, neww) > max_size:
        scale = max_size * 1.0 / max(newh, neww)
        newh = newh * scale
        neww = neww * scale
    return int(newh + 0.5), int(neww + 0.5)


def augmented_shapes(image_hw, min_sizes, max_size, flip):
    """-> [(new_h, new_w, flipped)] in the mapper's order."""
    out = []
    for size in min_sizes:
        nh, nw = resize_shortest_edge_shape(image_hw[0], image_hw[1], size, max_size)
        out.append((nh, nw, False))
        if flip:
            out.append((nh, nw, True))
    return out
