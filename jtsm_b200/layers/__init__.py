from .roi_align import ROIAlign, roi_align
from .roi_pool import RoIPool, roi_pool

__all__ = ["RoIPool", "roi_pool", "ROIAlign", "roi_align"]
