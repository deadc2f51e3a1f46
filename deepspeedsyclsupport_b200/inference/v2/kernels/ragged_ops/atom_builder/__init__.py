from .atom_builder import AtomBuilder  # noqa: F401
