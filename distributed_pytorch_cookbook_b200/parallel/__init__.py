from .ddp import DDPStrategy, GradReducer, plan_buckets  # noqa: F401
from .pipe import PipelineStage, PipeStrategy, partition_layers, schedule_actions, stage_layer_range  # noqa: F401
from .fsdp import FSDPStrategy, ShardGroup, plan_groups, plan_units  # noqa: F401
