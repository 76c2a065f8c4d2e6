import torch
import torch.nn.functional as F


class VaeImageProcessor:
    """Tensor-only stand-in for diffusers' VaeImageProcessor.preprocess: resize to (height, width) when needed, map [0,1] to
    [-1,1] when do_normalize (and the input is not already signed), binarize when asked."""

    def __init__(self, vae_scale_factor=8, do_normalize=True, do_binarize=False, do_convert_grayscale=False, **_):
        self.vae_scale_factor, self.do_normalize = vae_scale_factor, do_normalize
        self.do_binarize, self.do_convert_grayscale = do_binarize, do_convert_grayscale

    def preprocess(self, image, height=None, width=None):
        assert torch.is_tensor(image) and image.dim() == 4
        if self.do_convert_grayscale and image.shape[1] == 3:
            image = image.mean(dim=1, keepdim=True)
        if height is not None and tuple(image.shape[-2:]) != (height, width):
            image = F.interpolate(image, size=(height, width))
        if self.do_normalize and float(image.min()) >= 0:
            image = 2.0 * image - 1.0
        if self.do_binarize:
            image = (image >= 0.5).to(image.dtype)
        return image
