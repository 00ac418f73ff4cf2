// Pothos/Config.hpp -- stand-in: LoRa-SDR/ChirpGenerator.hpp includes it and uses nothing from it.
// TEST INFRASTRUCTURE ONLY (see Framework.hpp in this directory).
#pragma once
