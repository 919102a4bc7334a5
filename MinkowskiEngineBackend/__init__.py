"""Drop-in for the reference's compiled extension package
(`MinkowskiEngineBackend._C`, built from pybind/minkowski.cu)."""
