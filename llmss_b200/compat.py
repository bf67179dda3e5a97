"""Import-path compatibility with the reference package.

`import llmss_b200.compat` registers this framework's modules under the reference's names, so code
written against `jongwon-jay-lee/llmss` keeps working unchanged:

    from llmss.server.models.custom_modeling import MODEL_REGISTRY
    from llmss.server.models.utils.dist import initialize_torch_distributed
    from llmss.server.models.utils.hub import weight_files
    from llmss.server.models.utils.weights import Weights
    from llmss.server.models.utils.layers import TensorParallelRowLinear, ...

(reference: src/llmss/server/models/{custom_modeling,utils}/).  It refuses to shadow a real `llmss`
installation that is already imported.
"""
import sys
import types

from . import models as _models
from .utils import dist as _dist
from .utils import hub as _hub
from .utils import layers as _layers
from .utils import weights as _weights


def install(force: bool = False) -> None:
    if "llmss" in sys.modules and not getattr(sys.modules["llmss"], "__llmss_b200_compat__", False) and not force:
        raise RuntimeError("a real `llmss` package is already imported; refusing to shadow it (pass force=True)")

    def pkg(name):
        m = types.ModuleType(name)
        m.__path__ = []  # mark as package
        m.__llmss_b200_compat__ = True
        sys.modules[name] = m
        return m

    root, server, mods, utils = pkg("llmss"), pkg("llmss.server"), pkg("llmss.server.models"), pkg("llmss.server.models.utils")
    cm = pkg("llmss.server.models.custom_modeling")
    cm.MODEL_REGISTRY = _models.MODEL_REGISTRY
    for cls in ("GPTJForCausalLM", "GPTBigCodeForCausalLM", "GPT2LMHeadModel", "OPTForCausalLM", "LlamaForCausalLM"):
        setattr(cm, cls, getattr(_models, cls))
    root.server, server.models, mods.utils, mods.custom_modeling = server, mods, utils, cm
    for short, mod in (("dist", _dist), ("hub", _hub), ("layers", _layers), ("weights", _weights)):
        sys.modules[f"llmss.server.models.utils.{short}"] = mod
        setattr(utils, short, mod)


install()
