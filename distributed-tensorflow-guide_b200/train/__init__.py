"""``dtg.train`` -- the names the reference reaches through ``tf.train``."""
from .optimizers import (AdagradOptimizer, GradientDescentOptimizer, Optimizer, OutOfRangeError,
                         SyncReplicasOptimizer, get_global_step, get_or_create_global_step,
                         global_step_value)
from ..cluster import ClusterSpec, Server
from ..placement import replica_device_setter
from ..session.checkpoint import Saver, get_checkpoint_state, latest_checkpoint, load_checkpoint
from ..session.hooks import (CheckpointSaverHook, FeedFnHook, FinalOpsHook, GlobalStepWaiterHook,
                             HeartbeatHook, LoggingTensorHook, NanLossDuringTrainingError,
                             NanTensorHook, SessionRunArgs, SessionRunContext, SessionRunHook,
                             SessionRunValues, StepCounterHook, StopAtStepHook, SummarySaverHook,
                             SyncReplicasOptimizerHook)
from ..session.monitored import (ChiefSessionCreator, Coordinator, MonitoredSession,
                                 MonitoredTrainingSession, Scaffold, SessionCreator,
                                 SessionManager, Supervisor, WorkerSessionCreator)
from ..utils.trace import PhaseTracer, ProfilerHook
