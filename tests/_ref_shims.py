"""Stand-ins for the reference's missing third-party imports; they now live with the
hosting code in ``mdistiller_ddp_b200.refhost`` -- this module keeps the old test-side names."""

import mdistiller_ddp_b200  # noqa: F401
from mdistiller_ddp_b200.refhost import CfgNode, SummaryWriter, install_import_shims as install_shims  # noqa: F401
