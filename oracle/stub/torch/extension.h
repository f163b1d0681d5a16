// Empty stand-in for <torch/extension.h>: the reference's sampler headers include it but the
// HOST_DEVICE functions the oracle build uses need nothing from it. Test infrastructure only.
#pragma once
