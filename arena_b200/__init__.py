"""arena_b200 -- B200-native GPU resource accounting for `arena top node` /
`arena top job` (reference: mark1222/arena, cmd/arena/commands/{gpu,top_node,
top_job}.go).  The product is arena_b200/libarena_gpuacct.so (hand-written
sm_100a kernels behind a C ABI, include/*.h); this package is the thin host
mirror used by the tests, the bench and the multi-GPU driver."""

from ._lib import ARENA_NONE, ArenaError, init, init_devices, lib  # noqa: F401
from .snapshot import (DeviceSnapshot, HostSnapshot, MultiSnapshot, PackedDeviceSnapshot, PackedSnapshot,  # noqa: F401
                       PinnedBuffer)
from .top import (TopJobResult, TopNodeResult, format_job_list, format_top_job, format_top_node,  # noqa: F401
                  quantity_values, top_job, top_job_device, top_job_device_noalloc, top_node,
                  top_node_device, top_node_device_fused, top_node_device_posted, peer_exchange_collect,
                  top_node_packed, top_node_packed_device, top_node_multi, top_job_multi)
