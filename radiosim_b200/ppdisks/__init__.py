from .ppdisks import create_proto, generate_proto_set

__all__ = ["create_proto", "generate_proto_set"]
