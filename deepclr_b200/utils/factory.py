"""Construct a (transitive) subclass by its class name; same contract as the reference's deepclr/utils/factory.py:4-20."""
from typing import Any


def factory(cls: Any, name: str, *args: Any, **kwargs: Any) -> Any:
    stack = list(cls.__subclasses__())
    while stack:
        sub = stack.pop(0)
        if sub.__name__ == name:
            return sub(*args, **kwargs)
        stack = list(sub.__subclasses__()) + stack      # depth-first, declaration order
    raise NotImplementedError("Class '{}' not found as subclass of '{}'".format(name, cls.__name__))
