from .checkpoint import Saver, latest_checkpoint, get_checkpoint_state, load_checkpoint
from .hooks import *  # noqa: F401,F403
from .monitored import (ChiefSessionCreator, Coordinator, MonitoredSession,
                        MonitoredTrainingSession, Scaffold, Session, SessionCreator,
                        SessionManager, Supervisor, WorkerSessionCreator)
