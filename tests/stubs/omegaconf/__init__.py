"""Inert stand-in for `omegaconf` (absent from this image), imported at module level by vllm_omni/config/yaml_util.py.
The stage-YAML machinery that would call into it is not on the path the tests drive; anything that does call it
gets a placeholder back rather than a configuration."""


class _Anything:
    def __init__(self, name="omegaconf"):
        self._name = name

    def __call__(self, *a, **k):
        return _Anything(self._name + "()")

    def __getattr__(self, item):
        if item.startswith("__") and item.endswith("__"):
            raise AttributeError(item)
        return _Anything(f"{self._name}.{item}")

    def __iter__(self):
        return iter(())

    def __repr__(self):
        return f"<stub {self._name}>"


class DictConfig(dict):
    pass


class ListConfig(list):
    pass


OmegaConf = _Anything("OmegaConf")


def __getattr__(name):
    if name.startswith("__") and name.endswith("__"):
        raise AttributeError(name)
    return _Anything(f"omegaconf.{name}")
