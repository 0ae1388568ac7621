"""Stack B parameter manager: ParamShard + SingaServer/PMServer + SingaClient/PMClient configured by the
``Topology`` proto (server groups, neighbour groups, sync_interval)."""
from .param_shard import ParamShard
from .pm_client import PMClient, SingaClient
from .pm_server import PMServer, SingaServer
