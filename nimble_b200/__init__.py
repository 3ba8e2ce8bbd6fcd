"""nimble_b200: B200-native multi-chain MCMC back end behind NIMBLE's sampling API.

Host layer (this package, Python + ctypes because R is not available to the build) mirrors
nimbleCode / nimbleModel / configureMCMC / buildMCMC / compileNimble / runMCMC; the product is the
CUDA library nimble_b200/libnimb200.so behind the C ABI of include/nimb200.h."""
from .model import Model, ModelError, nimbleCode, nimbleModel
from .mcmc import (MCMCconf, CompiledMCMC, SampleMatrix, buildMCMC, check_conjugacy, compileNimble, configureMCMC,
                   nimbleOptions, runMCMC)
from .ess import effective_size, min_ess, samples_summary

__all__ = ["Model", "ModelError", "nimbleCode", "nimbleModel", "MCMCconf", "CompiledMCMC", "SampleMatrix",
           "buildMCMC", "check_conjugacy", "compileNimble", "configureMCMC", "runMCMC", "nimbleOptions", "effective_size", "min_ess",
           "samples_summary"]
