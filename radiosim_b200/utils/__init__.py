from .gauss import skewed_gauss, twodgaussian
from .noise import add_noise
from .utils import (_gen_date, _gen_vlba_obs_position, _save_mojave_bundle, cart2pol, create_grid, load_data,
                    pol2cart, relativistic_boosting, save_sky_distribution_bundle, zoom_on_source, zoom_out)

__all__ = ["add_noise", "cart2pol", "create_grid", "load_data", "pol2cart", "relativistic_boosting",
           "save_sky_distribution_bundle", "skewed_gauss", "twodgaussian", "zoom_on_source", "zoom_out"]
