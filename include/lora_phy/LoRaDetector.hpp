// lora_phy/LoRaDetector.hpp -- B200 build of the `lora_phy` API.
//
// The reference's detector (FFT + first-strict-maximum argmax over |X|^2 + power/fIndex read-outs, reference
// include/lora_phy/LoRaDetector.hpp:33-74) is fused into the demodulation kernels of this build
// (csrc/demod_fast.cu window_top / prepass_blocked_kernel); there is no host-side detector object.  The name is
// only declared so that translation units which mention the type in a pointer keep compiling.
#pragma once

template <typename T> class LoRaDetector;
