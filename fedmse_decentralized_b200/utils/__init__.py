from .similarity import (fusion_weights, jensenshannon, js_divergence, kde_self_log_density,
                         kl_divergence, scott_bandwidth, similarity_score)
from .timers import PhaseTimer

__all__ = ["similarity_score", "kl_divergence", "js_divergence", "kde_self_log_density",
           "jensenshannon", "scott_bandwidth", "fusion_weights", "PhaseTimer"]
