"""Model zoo used by the benchmarks: every builder returns a flat ``nn.Sequential``
whose layer names (and therefore ``state_dict`` keys) equal those of the
reference's ``benchmarks/models``."""
from torchgpipe_b200.models.amoebanet import amoebanetd
from torchgpipe_b200.models.resnet import resnet101
from torchgpipe_b200.models.unet import unet
from torchgpipe_b200.models.flatten import flatten_sequential
from torchgpipe_b200.models.vgg import resnet50_coarse, resnet101_coarse, vgg16

__all__ = ['amoebanetd', 'resnet101', 'unet', 'flatten_sequential', 'vgg16', 'resnet101_coarse',
           'resnet50_coarse']
