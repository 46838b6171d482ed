// Stand-in for a header of the hyteg_operators submodule, which is absent from the reference tree (.SUBMODULES.json: missing).
// src/mixed_operator/VectorLaplaceOperator.hpp only names the type in a typedef and an extern template declaration.
#pragma once
namespace hyteg { namespace operatorgeneration { class P2PlusBubbleElementwiseDiffusion; } }
