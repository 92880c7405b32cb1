from .communicator import Communicator, DistCommunicator, InProcessCommunicator  # noqa: F401
from .coordinator import Coordinator, DistributedCoordinator, InProcessCoordinator, PipelineTimeout  # noqa: F401
from .protocol import CommandType, Endpoint, Job, LoadTracker, Message, StageConfig  # noqa: F401
from .stage import NetworkStageWorker, PipelineStage  # noqa: F401
from .engine import PipelineEngine, PipelineLinks, build_schedule, make_pipeline_groups  # noqa: F401
from .tcp import PipelineStageLost, TcpCommunicator, TcpCoordinator, TcpStageWorker  # noqa: F401
