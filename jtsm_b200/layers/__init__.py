from .roi_pool import RoIPool, roi_pool

__all__ = ["RoIPool", "roi_pool"]
