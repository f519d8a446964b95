// peer.h — the per-period exchange of the path-sharded valuation over peer memory (NVLink / NVSwitch), SURVEY §8(e).
//
// Every rank (one lsmc_ctx per GPU; ranks are processes or host threads) owns one exchange buffer in its own HBM:
//     slots   [2 parities][world][kPeerSlotDoubles]      what rank r contributed in the current / previous exchange
//     counter one 64-bit arrival counter
// and holds device pointers to the buffers of all ranks (cudaIpc-mapped across processes, plain pointers inside one
// process).  An exchange is: write my block into slot [parity][my rank] of EVERY rank's buffer with plain peer stores,
// fence at system scope, take a ticket on this GPU; the CTA with the last ticket adds the rank's CTA count to every rank's
// counter (world remote atomics per rank and exchange, not CTAs x world); spin until my own counter has reached the cumulative
// target (every CTA of every rank has arrived); sum the world blocks of my buffer in rank order.  All ranks sum the same
// values in the same order, so the result — and the regression coefficients solved from it — is bit-identical on every
// rank and does not depend on a collective library's ring order.  Two parities suffice: nobody can start exchange e+2
// before everybody has finished reading exchange e (a rank must arrive at e+1 for anyone to leave e+1).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "kernels.h"

namespace lsmc {

constexpr int kMaxPeers = 8;
constexpr int kPeerSlotDoubles = 24 * 1024;   // per (parity, rank): >= the largest normal-equation block (K = 64, 26 + 8 column blocks)

// Device-visible view of the communicator, passed to the kernels by value.
struct PeerView {
    int rank, world;
    double* slots[kMaxPeers];                 // rank r's slot array
    unsigned long long* counter[kMaxPeers];   // rank r's arrival counter
    unsigned long long* ticket;               // this rank's own CTA ticket (device scope): the last CTA of an exchange signals the peers
};

struct TailParams {
    SolveParams solve;          // solve.reduced is not used: the reduced block is read from the slots
    const double* partial;      // [n_parts][n_elem] per-CTA partials of the sweep
    int n_parts, n_elem;
    int do_solve;               // 0: only reduce + exchange (period 0: the backward NPV)
    double* total_out;          // [n_out] the first elements of the exchanged block, written by CTA 0 (host reads them)
    int n_out;
    PeerView peers;
    int parity;                 // this exchange's slot parity
    unsigned long long target;  // cumulative arrivals at this rank's counter after this exchange
    unsigned long long local_target;   // cumulative tickets of this rank's own CTAs after this exchange
    int* error_flag;            // set to kPeerTimeout when a spin times out
    long long* trace;           // development (LSMC_TAIL_TRACE): globaltimer stamps of CTA 0, [8]; NULL otherwise
    int pdl;                    // host side: launch with programmatic stream serialisation
};
constexpr int kPeerTimeout = 9;

int tail_blocks(int n_elem, int G, int num_sms);
void launch_tail(const TailParams& tp, int blocks, cudaStream_t stream);

// Small collectives through the same buffers: in-place all-reduce (sum in rank order) or broadcast from `root` of
// buf[0, count), count <= kPeerSlotDoubles.  One CTA.
struct PeerSmallParams {
    double* buf;
    int count, root;            // root < 0: all-reduce; root >= 0: broadcast from that rank
    PeerView peers;
    int parity;
    unsigned long long target, local_target;
    int* error_flag;
};
void launch_peer_small(const PeerSmallParams& pp, cudaStream_t stream);

}  // namespace lsmc
