/* TEST INFRASTRUCTURE ONLY: empty stand-in for ggml-quants.h (see ggml-bitnet.h in this directory). */
#pragma once
