"""Plugin loading, same contract as gaudi/plugin.py:69-122: plugins are named by dotted
module path in the job description; each module exposes ``enable(**kwargs)``."""
from __future__ import annotations

import sys
from collections import OrderedDict
from importlib import import_module


def import_plugins(*pluginlist):
    """gaudi/plugin.py:69-92."""
    plugins = []
    for plugin in pluginlist:
        module = plugin.module
        if module not in sys.modules:
            import_module(module)
        plugins.append(sys.modules[module])
    return plugins


def load_plugins(plugins, container=None, **kwargs):
    """gaudi/plugin.py:95-122: ``container[p.name] = sys.modules[p.module].enable(**kwargs, **p)``."""
    if container is None:
        container = OrderedDict()
    for p in plugins:
        if p.name not in container:
            plugin_kwargs = kwargs.copy()
            plugin_kwargs.update(p)
            if p.module not in sys.modules:
                import_module(p.module)
            container[p.name] = sys.modules[p.module].enable(**plugin_kwargs)
    return container
