"""Head modules of the WSL ROI heads with the reference's class names and signatures on the sm_100a kernels."""
from .box_head import DiscriminativeAdaptionNeck
from .fast_rcnn_oicr import OICROutputLayers
from .fast_rcnn_wsddn import WSDDNOutputLayers
from .roi_heads_jtsm import (CMILROIHeadsB200, ContextLocNetROIHeadsB200, JTSMROIHeadsB200, PCLROIHeadsB200,
                             WSDDNROIHeadsB200, register_all)
from .roi_heads_oicr import OICRROIHeadsB200, get_image_level_gt, register

__all__ = ["DiscriminativeAdaptionNeck", "OICROutputLayers", "WSDDNOutputLayers", "OICRROIHeadsB200",
           "WSDDNROIHeadsB200", "ContextLocNetROIHeadsB200", "CMILROIHeadsB200", "PCLROIHeadsB200", "JTSMROIHeadsB200", "get_image_level_gt", "register", "register_all"]
