"""``tf.train``-shaped namespace: cluster/server, placement, optimizers, sessions, hooks, saver."""
from ..cluster import ClusterSpec  # noqa: F401
from ..framework.graph import (create_global_step, get_global_step,  # noqa: F401
                               get_or_create_global_step)
from ..placement import (FlatSlicedStrategy, GreedyLoadBalancingStrategy,  # noqa: F401
                         RoundRobinStrategy, byte_size_load_fn, replica_device_setter)
from .hooks import (CheckpointSaverHook, CheckpointSaverListener, DataPlaneStatsHook,  # noqa: F401
                    FaultInjectionHook,
                    FeedFnHook, FinalOpsHook, GlobalStepWaiterHook, LoggingTensorHook,
                    NanLossDuringTrainingError, NanTensorHook, SecondOrStepTimer, SessionRunArgs,
                    SessionRunContext, SessionRunHook, SessionRunValues, StepCounterHook,
                    StopAtStepHook, SummarySaverHook)
from .optimizers import (AdagradOptimizer, AdamOptimizer, GradientDescentOptimizer,  # noqa: F401
                         MomentumOptimizer, Optimizer, OptimizerSpec)
from .saver import (CheckpointState, Saver, checkpoint_exists, get_checkpoint_state,  # noqa: F401
                    latest_checkpoint, update_checkpoint_state)
from .server import LocalServer, Server  # noqa: F401
from .session import (AbortedError, ChiefSessionCreator, ConfigProto, Coordinator,  # noqa: F401
                      MonitoredSession, MonitoredTrainingSession, OutOfRangeError, Scaffold,
                      Session, SessionManager, SingularMonitoredSession, UnavailableError,
                      WorkerSessionCreator, global_variables_initializer,
                      report_uninitialized_variables)
from .supervisor import Supervisor  # noqa: F401
from .sync_replicas import SyncReplicasOptimizer  # noqa: F401
from ..utils.events import summary_iterator  # noqa: F401
