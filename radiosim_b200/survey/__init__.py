from .survey import create_survey

__all__ = ["create_survey"]
