from .mojave import (add_swirl, create_mojave, gen_compact, gen_jet, gen_one_jet, gen_shocks, gen_two_jet,
                     swirl_image)

__all__ = ["add_swirl", "create_mojave", "gen_compact", "gen_jet", "gen_one_jet", "gen_shocks", "gen_two_jet",
           "swirl_image"]
