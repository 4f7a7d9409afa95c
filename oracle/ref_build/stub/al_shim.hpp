// ORACLE BUILD ONLY. The reference's src/core/imports/mpi.cpp names Aluminum-only symbols inside
// discarded `if constexpr` branches of TaggedSend/TaggedRecv (mpi.cpp:598-640, 891-933), which a
// CPU-only build (no Aluminum) cannot parse.  IsAluminumSupported is std::false_type there
// (include/El/core/imports/aluminum.hpp:41-47), so the branches are never taken; this header gives
// them something to parse against.  Force-included for that one translation unit.
#pragma once
#include <El-lite.hpp>
namespace El {
template <typename T, Device D, Collective C> using BestBackend = void;
}
namespace Al {
template <typename Backend, typename... Args> void Send(Args&&...) {}
template <typename Backend, typename... Args> void Recv(Args&&...) {}
}
