// handoff.cuh -- stage-to-stage activation handoff rings (placeholder until the
// peer-memory implementation lands; every entry point fails loudly).
#pragma once
#include "common.cuh"

extern "C" {
int b2p_ring_create(b2p_ring** out, int, int, size_t) { if (out) *out = nullptr; return b2p::fail(B2P_ESTATE, "ring: not built yet"); }
int b2p_ring_export(b2p_ring*, void*) { return b2p::fail(B2P_ESTATE, "ring: not built yet"); }
int b2p_ring_open(b2p_ring** out, const void*, int) { if (out) *out = nullptr; return b2p::fail(B2P_ESTATE, "ring: not built yet"); }
int b2p_ring_open_local(b2p_ring** out, b2p_ring*, int) { if (out) *out = nullptr; return b2p::fail(B2P_ESTATE, "ring: not built yet"); }
int b2p_ring_destroy(b2p_ring*) { return 0; }
int b2p_ring_send(b2p_ring*, const void*, size_t, uint64_t, void*) { return b2p::fail(B2P_ESTATE, "ring: not built yet"); }
int b2p_ring_recv(b2p_ring*, uint64_t, void**, void*) { return b2p::fail(B2P_ESTATE, "ring: not built yet"); }
int b2p_ring_release(b2p_ring*, uint64_t, void*) { return b2p::fail(B2P_ESTATE, "ring: not built yet"); }
}
