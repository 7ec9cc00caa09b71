from .mojave import create_mojave, gen_compact, gen_one_jet, gen_two_jet

__all__ = ["create_mojave", "gen_compact", "gen_one_jet", "gen_two_jet"]
