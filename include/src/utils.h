// Intentionally empty.  The reference's src/utils.h is a fully commented-out copy of device
// helpers (src/utils.h:1-634) that examples/cpp/testFlashDecoder.cc:4 still includes.
#pragma once
