"""B200-native parameter-server training framework with the capabilities of the
Distributed-TensorFlow-Guide examples (see SURVEY.md / DESIGN.md).

Public surface (``import dtg_b200 as dtg``):

* cluster / roles : ``ClusterSpec``, ``Server``, ``ConfigProto``, ``GPUOptions``
* variables       : ``Variable``, ``GraphKeys``, ``device``, ``replica_device_setter``,
  collections and init ops, lazy ``Tensor`` / ``Op`` nodes
* optimizers      : ``dtg.train.GradientDescentOptimizer`` / ``AdagradOptimizer`` /
  ``SyncReplicasOptimizer``
* lifecycle       : ``dtg.train.MonitoredTrainingSession`` / ``Supervisor`` / ``Scaffold`` /
  hooks / ``Saver``
* schemes         : ``dtg.parallel`` engines (Hogwild, DOWNPOUR, DOWNPOUR-Easy, ADAG,
  SSGD, per-worker-LR SSGD, SDAG, multi-GPU async and tower averaging)
* kernels         : ``dtg.ops`` (sm_100a CUDA extension)
"""
from .cluster import (ClusterSpec, ConfigProto, DeviceAttributes, DeviceSpec, GPUOptions,
                      Server, canonical_device)
from .framework import (GraphKeys, Graph, Op, Tensor, Variable, abs, add_to_collection, assign,
                        assign_add, assign_sub, constant, control_dependencies, device,
                        get_collection, get_collection_ref, get_default_graph, get_variable,
                        global_variables, global_variables_initializer, group, local_variables,
                        local_variables_initializer, name_scope, no_op, op, reduce_mean,
                        reduce_sum, report_uninitialized_variables, reset_default_graph,
                        set_random_seed, square, stack, tensor, trainable_variables,
                        truncated_normal, variables_initializer, zeros)
from .placement import replica_device_setter
from .session.monitored import Session
from . import runtime
from . import train
from . import errors

__version__ = "0.1.0"


def reset():
    """Fresh graph + runtime (tests, notebooks)."""
    srv = runtime.current_server(required=False)
    if srv is not None:
        srv.stop()
    runtime.reset()
    return reset_default_graph()
