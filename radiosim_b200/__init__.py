"""radiosim_b200 -- B200-native (sm_100a CUDA) drop-in for radiosim's image synthesis.

Mirrors the reference's import paths for the hot path only:

    radiosim_b200.jets.create_jet          <- radiosim.jets.create_jet
    radiosim_b200.survey.create_survey     <- radiosim.survey.create_survey
    radiosim_b200.mojave.create_mojave     <- radiosim.mojave.create_mojave
    radiosim_b200.ppdisks.generate_proto_set / create_proto
    radiosim_b200.utils.{twodgaussian, skewed_gauss, zoom_on_source, zoom_out, ...}

All pixel work runs in ``librsb200.so`` (see ``include/radiosim_b200.h``); there is
no CPU fallback.
"""

__version__ = "0.1.0"
