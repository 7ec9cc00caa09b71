from .jet import apply_train_type, component_from_list, create_jet

__all__ = ["apply_train_type", "component_from_list", "create_jet"]
