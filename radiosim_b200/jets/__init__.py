from .jet import apply_train_type, component_from_list, create_jet
from .packer import draw_from_mojave_dist, get_start_amp

__all__ = ["apply_train_type", "component_from_list", "create_jet", "draw_from_mojave_dist", "get_start_amp"]
