from .moi_pool import MOIPool, moi_pool
from .pcl_loss import PCLLoss, pcl_loss
from .roi_align import ROIAlign, roi_align
from .roi_label import ROILabel
from .roi_loop_pool import ROILoopPool, roi_loop_pool
from .roi_merge import ROIMerge, roi_merge_padded
from .roi_pool import RoIPool, roi_pool

__all__ = ["RoIPool", "roi_pool", "ROIAlign", "roi_align", "ROILoopPool", "roi_loop_pool", "MOIPool", "moi_pool",
           "PCLLoss", "pcl_loss", "ROIMerge", "roi_merge_padded", "ROILabel"]
