from .gauss import skewed_gauss, twodgaussian
from .utils import (create_grid, pol2cart, relativistic_boosting, save_sky_distribution_bundle,
                    zoom_on_source, zoom_out)

__all__ = ["create_grid", "pol2cart", "relativistic_boosting", "save_sky_distribution_bundle",
           "skewed_gauss", "twodgaussian", "zoom_on_source", "zoom_out"]
