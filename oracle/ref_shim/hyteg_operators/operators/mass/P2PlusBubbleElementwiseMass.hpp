// Stand-in for a header of the absent hyteg_operators submodule (see P2PlusBubbleElementwiseDiffusion.hpp next to it).
#pragma once
namespace hyteg { namespace operatorgeneration { class P2PlusBubbleElementwiseMass; } }
