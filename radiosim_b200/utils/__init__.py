from .gauss import skewed_gauss, twodgaussian
from .noise import add_noise
from .utils import (create_grid, pol2cart, relativistic_boosting, save_sky_distribution_bundle,
                    zoom_on_source, zoom_out)

__all__ = ["add_noise", "create_grid", "pol2cart", "relativistic_boosting", "save_sky_distribution_bundle",
           "skewed_gauss", "twodgaussian", "zoom_on_source", "zoom_out"]
