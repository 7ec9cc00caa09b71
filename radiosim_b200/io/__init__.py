from .config import Config, ConfigError
