// srb200.cu -- C ABI (include/srb200.h) over the sm_100a kernels.
//
// One SrbContext = one GPU, one CUDA stream, one framebuffer (or band of it).
// srb_draw* only records commands; srb_end_frame uploads the draw table and
// launches the whole pipeline asynchronously:
//   frustum -> vertex -> setup(+count) -> scan -> fill -> [sort ->] tile raster.
// There is no CPU fallback: without a CUDA device srb_create fails.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "srb_raster.cuh"
#include "srb_physics.cuh"

using namespace srb;

namespace {

thread_local std::string g_createError;

struct MeshRes { float4* vtx = nullptr; uint16_t* idx = nullptr; float4* blockBoxes = nullptr; int2* blockVRange = nullptr; int nV = 0, nI = 0; bool opaque = true, alive = false; };
struct TexRes { uchar4* px = nullptr; int w = 0, h = 0; bool opaque = true, alive = false; };

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 4 + 64;
        cudaError_t e = cudaMalloc((void**)&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct QueuedDraw { SrbDrawCmd cmd; int64_t order; };

}  // namespace

struct SrbContext {
    int device = 0, smCount = 148;
    int W = 0, H = 0, tilesX = 0, tilesY = 0;
    int band0 = 0, band1 = 0;
    bool bandSet = false;
    int pipeline = SRB_PIPELINE_AUTO;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float4* color = nullptr;
    float* depth = nullptr;
    float4* colorTarget = nullptr;    // peer buffer opened from an IPC handle (band gather), or null
    uint32_t* color8 = nullptr;       // RGBA8 gather frame owned by this context (rank 0 of an RGBA8 band gather)
    uint32_t* color8Target = nullptr; // a peer's RGBA8 gather frame (band ranks)
    // present path: packed RGB float copies of finished frames, double buffered so that the D2H copy of
    // frame k (copy stream) overlaps the rendering of frame k+1 (pipeline stream)
    cudaStream_t copyStream = nullptr;
    DevBuf<float> dPresent[2];
    static constexpr int bigCountCtas = 1;   // CTAs per SM of the large-triangle count pass (measured: 1 / 2 / 4)
    static constexpr int fillCtas = 4;       // CTAs per SM and role of the fill pass
    // (The count pass waits on atomic round trips, not on issue slots: with 4 CTAs per SM it is a few us shorter alone
    // -- 45.5 vs 50 us for setup+count -- but holds four times the registers while the previous frame's tile kernel
    // wants them; pipelined, city4k gains about 1 % with 1.)
    DevBuf<float4> dRayScratch;       // srb_raycast: directions, hits, keys, rays
    cudaEvent_t presentReady[2] = {}, presentCopied[2] = {};
    bool presentBusy[2] = {false, false};
    int presentSlot = 0;
    std::vector<MeshRes> meshes;      // handle = index + 1
    std::vector<TexRes> textures;
    std::mutex mu;
    SrbUniforms uniforms;
    // frame recording
    bool inFrame = false, haveFrame = false, framePending = false;
    float clear[4] = {0, 0, 0, 1};
    uint32_t clearFlags = 0;
    std::vector<QueuedDraw> queue;
    int64_t callCounter = 0;
    // resident frame
    FrameParams fp;
    bool useVis = false;
    int visMode = 0;
    int nVBlocks = 0, nTBlocks = 0;
    DevBuf<uint32_t> dBinScratch;
    // peer-memory frame barrier (multi-GPU bands, one rank per GPU)
    uint32_t* barRow = nullptr;                 // this rank's row of flags (exported)
    BarrierRows barRows = {};                   // every rank's row: own + peers' opened through IPC handles
    int barRank = -1, barWorld = 0;
    uint32_t barCount = 0;                      // barriers issued so far
    uint32_t* barTimedOut = nullptr;            // pinned, mapped
    cudaStream_t barStream = nullptr;           // slack >= 1: the barrier kernel runs beside the raster stream
    cudaEvent_t barFrame = nullptr, barDone[2] = {};
    int barPending = -1;                        // barDone[] index of the last side-stream barrier, -1: none
#ifdef SRB_TIMELINE
    unsigned long long* dTimeline = nullptr;    // diagnostic builds: globaltimer at four points of every frame
    uint32_t timelineFrame = 0;
#endif
    unsigned long long* dSliceKeys = nullptr;   // key slices of split bins (kMaxExtra x 256), all zero between tile kernels
    uint32_t* dSliceDone = nullptr;             // per slice: parts merged so far, zero between tile kernels
    // Bins of >= splitMin entries are rasterized by several CTAs of splitPart entries each (srb_set_bin_split).
    // A full frame is bound by the sum of its tiles, not by the longest one, and splitting costs a merge, so it
    // starts late there; a band's floor IS its heaviest tile.
    uint32_t splitMinFull = 0, splitMinBand = 256, splitPart = 128;   // (with hi-Z: profiles/r02z_band_probe_split_under_hiz.txt)
    // Frame pipelining: the geometry stages of the next frames run on their own streams while the tile kernel of
    // frame N is still running on `stream`.  Everything a frame's stages write exists once per SET; frames take the
    // sets in turn, each set has its own geometry stream, and a set is reused when the tile kernel of its previous
    // frame is done.  With two sets the geometry of frame N+1 overlaps the tile kernel of frame N; with more, the
    // geometry chains of consecutive frames -- dependent, latency-bound launches that leave most of the machine idle --
    // also overlap each other, which is what a band context with a short tile kernel needs (srb_set_pipeline_depth).
    static constexpr int kMaxSets = 4;
    static constexpr int nSets = kMaxSets;    // every set is allocated; `depth` says how many take turns
    bool useGraphs = true;                    // srb_set_graphs
    int pipeDepth = 0;                        // 0 = chosen per frame (launch_frame), else 2..4
    struct FrameBufs {
        cudaStream_t geomStream = nullptr;
        cudaEvent_t geomDone = nullptr, rasterDone = nullptr, geomStarted = nullptr;
        bool rasterRecorded = false, startedValid = false;
        // The set's geometry chain (one memset + 6..8 kernels) as an instantiated CUDA graph: a frame that repeats the
        // shape of the set's previous frame (same parameters, grids and buffers -- graphKey) costs the host one launch
        // instead of nine.  Captured the second time a shape is seen, dropped when it changes.
        cudaGraphExec_t graphExec = nullptr;
        std::vector<unsigned char> graphKey, lastKey;
        int graphLaunches = 0;
        DevBuf<int2> dVBlocks, dTBlocks;  // block tables of the vertex / setup kernels: (draw, first element) of every run of
                                          // kGeomBlock vertices / triangles of a VISIBLE draw, compacted per frame by the frustum kernel
        DevBuf<float4> dClip;             // clip positions (vertex stage -> setup)
        DevBuf<TriRef> dRefs;             // accepted triangles (setup -> fill)
        DevBuf<BigEntry> dBigQ;           // triangles that span more than kSmallTiles tiles
        DevBuf<uint32_t> dCntTiles;       // [64 words: Counters][per-tile counts][per-tile large-triangle flags]: one memset per frame
        DevBuf<DrawRec> dDraws;           // this set's copy of the draw table (the frustum kernel writes into it)
        uint64_t tableSerial = 0;         // which frame's table it holds
        DevBuf<float4> dWn;               // world normals (vertex stage -> shading)
        DevBuf<float4> dRecsR, dRecsS;    // split setup records (RecStore)
        DevBuf<ClipAttr> dClipAttrs;
        DevBuf<uint32_t> dTileStart, dTileCursor, dBins;
        DevBuf<uint4> dOrder;             // tile launch order of the visibility kernel
        Counters* dCnt = nullptr;         // = dCntTiles.p
        uint32_t* dTileCount = nullptr;   // = dCntTiles.p + 64
        uint32_t* dTileHasBig = nullptr;  // = dTileCount + tiles
        Counters* hCnt = nullptr;         // pinned; written by the scan kernel itself (every counter is final by then)
        cudaEvent_t cntReady = nullptr;   // ... and this event follows the copy: the host learns about a capacity
                                          // overflow when the frame's GEOMETRY is done, not its rasterization
    } fb[kMaxSets];
    int slot = 0;                         // set used by the most recently launched frame
    cudaStream_t lastGeomStream = nullptr;
    uint64_t frameSerial = 0;             // bumped by every srb_end_frame: the draw table in hDraws[tableSlot]
    int tableSlot = 0;
    DrawRec* hDraws[2] = {nullptr, nullptr};   // pinned staging of the draw table, two alternating buffers
    size_t hDrawsCap[2] = {0, 0};
    cudaEvent_t stagingDone[2] = {};  // the H2D copy out of hDraws[i] has finished
    bool stagingBusy[2] = {false, false};
    int stagingSlot = 0;
    SrbStats stats;
    bool profiling = false;
    cudaEvent_t stageEv[SRB_N_STAGES + 1] = {};
    bool stageRan[SRB_N_STAGES] = {};
    float stageMs[SRB_N_STAGES] = {};
    std::string err;
};

namespace {

void close_color_target(SrbContext* c) {
    if (c->colorTarget) { cudaIpcCloseMemHandle(c->colorTarget); c->colorTarget = nullptr; }
}
void barrier_close(SrbContext* c) {
    if (c->barStream) cudaStreamSynchronize(c->barStream);
    c->barPending = -1;
    for (int r = 0; r < c->barWorld; ++r)
        if (r != c->barRank && c->barRows.row[r]) cudaIpcCloseMemHandle(c->barRows.row[r]);
    c->barRows = BarrierRows{};
    c->barRank = -1; c->barWorld = 0;
}
void close_color8_target(SrbContext* c) {
    if (c->color8Target) { cudaIpcCloseMemHandle(c->color8Target); c->color8Target = nullptr; }
}

int fail(SrbContext* c, int code, const std::string& msg) {
    if (c) c->err = msg; else g_createError = msg;
    return code;
}
#define CK(c, call)                                                                           \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return fail((c), SRB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

void default_uniforms(SrbUniforms& u) {
    std::memset(&u, 0, sizeof(u));
    // Renderer.cs:39-44 defaults; EulerToDirection(-45,-45,0) is supplied by the caller normally.
    u.light_dir[0] = 0.5f; u.light_dir[1] = -0.70710677f; u.light_dir[2] = -0.5f;
    for (int i = 0; i < 4; ++i) u.light_color[i] = 1.0f;
    u.fog_color[0] = 1.0f; u.fog_color[1] = 0.62f; u.fog_color[2] = 0.5f; u.fog_color[3] = 1.0f;
    u.fog_start = 1.0f; u.fog_end = 25.0f;
    u.near_clip = 0.1f;
    u.numerics_fma = 1;
}

int alloc_framebuffer(SrbContext* c, int w, int h) {
    if (c->color) cudaFree(c->color);
    if (c->depth) cudaFree(c->depth);
    if (c->color8) cudaFree(c->color8);      // a resize drops the RGBA8 gather frame: export it again
    c->color = nullptr; c->depth = nullptr; c->color8 = nullptr;
    CK(c, cudaMalloc((void**)&c->color, (size_t)w * h * sizeof(float4)));
    CK(c, cudaMalloc((void**)&c->depth, (size_t)w * h * sizeof(float)));
    CK(c, cudaMemsetAsync(c->color, 0, (size_t)w * h * sizeof(float4), c->stream));
    CK(c, cudaMemsetAsync(c->depth, 0, (size_t)w * h * sizeof(float), c->stream));
    c->W = w; c->H = h;
    c->tilesX = (w + kTile - 1) / kTile;
    c->tilesY = (h + kTile - 1) / kTile;
    if (!c->bandSet) { c->band0 = 0; c->band1 = c->tilesY; }
    c->band0 = std::min(c->band0, c->tilesY);
    c->band1 = std::min(c->band1, c->tilesY);
    c->haveFrame = false;
    return SRB_OK;
}

// Frame-dependent capacities are the smallest over the sets in use.
void refresh_caps(SrbContext* c) {
    size_t pairs = ~(size_t)0, clips = ~(size_t)0, big = ~(size_t)0;
    for (int i = 0; i < c->nSets; ++i) {
        pairs = std::min(pairs, c->fb[i].dBins.cap);
        clips = std::min(clips, c->fb[i].dClipAttrs.cap);
        big = std::min(big, c->fb[i].dBigQ.cap);
    }
    c->fp.pairCap = (uint32_t)std::min<size_t>(pairs, 0xffffffffu);
    c->fp.clipCap = (uint32_t)std::min<size_t>(clips, 0xffffffffu);
    c->fp.bigCap = (uint32_t)std::min<size_t>(big, 0xffffffffu);
}

#ifdef SRB_TIMELINE
// Diagnostic builds only (tools/timeline.py): where a frame's geometry chain and tile kernel sit in time.
constexpr uint32_t kTimelineFrames = 4096;
__global__ void stamp_kernel(unsigned long long* out) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    *out = t;
}
#define SRB_STAMP(point, stream_) \
    stamp_kernel<<<1, 1, 0, stream_>>>(c->dTimeline + 4 * (size_t)(c->timelineFrame % kTimelineFrames) + (point))
#else
#define SRB_STAMP(point, stream_) ((void)0)
#endif

// Launch the whole pipeline for the resident draw list.
int launch_frame(SrbContext* c) {
    FrameParams& fp = c->fp;
    const SrbUniforms U = c->uniforms;
    const int nTiles = fp.tilesX * fp.tilesY;
    const int bandTiles = (fp.bandRow1 - fp.bandRow0) * fp.tilesX;
    // A context that renders the whole frame is bound by its tile kernel, and the geometry of the next frame beside it
    // should be lean (few, light CTAs) rather than quick; a band rank's tile kernel is shorter than its geometry chain,
    // so there the chain's duration is the period and the quicker variants of the binning passes are used.
    const bool fullFrame = fp.bandRow0 == 0 && fp.bandRow1 == fp.tilesY;
    int launches = 0;
    int stage = 0;
    const uint32_t prevRefs = c->fb[c->slot].hCnt->nRefs, prevBig = c->fb[c->slot].hCnt->nBig;   // last launched frame (heuristics only)
    // Grids of the vertex and setup kernels: the live block counts of the last launched frame plus a margin (their
    // stride loops cover whatever the frustum kernel of this frame really leaves alive); every block when unknown.
    const uint32_t prevV = c->fb[c->slot].hCnt->nVLive, prevT = c->fb[c->slot].hCnt->nTLive;
    // Never fewer than two full waves' worth of CTAs (when the scene has that many blocks): a frame with far more live
    // blocks than the one before it (alternating passes, a camera cut) then still runs at full occupancy.
    // (Measured, no effect, round 2: grids of the vertex / setup / count / fill kernels sized tightly from the last
    // frame's counters instead of these floors and whole waves -- band contexts +-1 %, full frame unchanged:
    // profiles/r02v_band_probe_lean_grids_experiment.txt.  CTAs that find nothing to do retire too quickly to matter.)
    const int floorCtas = c->smCount * 16;
    // (Rounded up to a multiple of 256: a frame that repeats keeps its grids, and with them its graph.)
    auto quant = [](uint32_t n) { return (int)((n + n / 8 + 16 + 255u) & ~255u); };
    const unsigned gridV = (unsigned)std::max(1, prevV ? std::min<int>(c->nVBlocks, std::max<int>(floorCtas, quant(prevV))) : c->nVBlocks);
    const unsigned gridT = (unsigned)std::max(1, prevT ? std::min<int>(c->nTBlocks, std::max<int>(floorCtas, quant(prevT))) : c->nTBlocks);
    // Frames in flight.  A frame whose tile kernel outlasts its geometry chain gains nothing from more than two sets
    // (measured, city4k: 5,950 frames/s with 2, 5,800 with 4 -- more geometry kernels beside the tile kernel only
    // take SMs from it); a frame with a short tile kernel is bound by the LATENCY of its chain of geometry launches,
    // and overlapping those chains is all there is to gain (gun1080: 17,600 -> 29,300; one tile row of city4k:
    // 14,400 -> 21,500).  The tile kernel's time is estimated from the last frame's counters.
    int depthNow = c->pipeDepth;
    if (depthNow == 0) {
        const float estTileUs = 0.9e-3f * (float)bandTiles + 0.26e-3f * (float)c->fb[c->slot].hCnt->nPairs;
        depthNow = estTileUs < 70.0f ? 4 : 2;
    }
    const int sPrev = c->slot;
    const int s = (c->slot + 1) % depthNow;
    c->slot = s;
    SrbContext::FrameBufs& B = c->fb[s];
    // geometry + binning stream; the tile kernel stays on c->stream.  Stage profiling runs everything on
    // c->stream: the per-stage events then bracket kernels only, not cross-stream hand-overs.
    cudaStream_t gs = c->profiling ? c->stream : B.geomStream;
    auto mark = [&](int next) {       // close stages [stage, next) at this point of the pipeline
        if (!c->profiling) return;
        for (; stage < next; ++stage) cudaEventRecord(c->stageEv[stage + 1], stage + 1 <= 5 ? gs : c->stream);
    };
    if (B.rasterRecorded) CK(c, cudaStreamWaitEvent(gs, B.rasterDone, 0));   // this set's previous frame
    // Geometry chains start in frame order (a later frame's short kernels must not take the SMs from an earlier
    // frame's): behind the previous frame's frustum kernel, not behind its whole chain.
    if (c->lastGeomStream && c->lastGeomStream != gs && c->fb[sPrev].startedValid)
        CK(c, cudaStreamWaitEvent(gs, c->fb[sPrev].geomStarted, 0));
    c->lastGeomStream = gs;
    CK(c, cudaEventRecord(c->ev0, gs));
    SRB_STAMP(0, gs);
    if (B.tableSerial != c->frameSerial && fp.nDraws > 0) {
        // Every set has its own device copy of the draw table, uploaded on the geometry stream the first time the
        // set renders a new frame: a new frame's geometry then does not wait for the previous frame's tile kernel.
        CK(c, cudaMemcpyAsync(B.dDraws.p, c->hDraws[c->tableSlot], (size_t)fp.nDraws * sizeof(DrawRec), cudaMemcpyHostToDevice, gs));
        CK(c, cudaEventRecord(c->stagingDone[c->tableSlot], gs));
        c->stagingBusy[c->tableSlot] = true;
    }
    B.tableSerial = c->frameSerial;
    const uint32_t splitMin = c->useVis ? (fullFrame ? c->splitMinFull : c->splitMinBand) : 0u;
    // (Measured without effect, round 2: the fill grid sized from the triangles the last frame accepted instead of
    // the scene's total, and 1 / 2 instead of 4 CTAs per SM for a band's large-triangle count pass -- band contexts
    // and full frames within 1 %.  profiles/r02zw_band_probe_lean_fill_grids.txt)
    const unsigned fillGrid = (unsigned)std::min<size_t>((2 * (size_t)fp.totalT + 255) / 256, (size_t)c->smCount * 8);
    const unsigned wave = (unsigned)c->smCount * c->fillCtas;
    // Fill: one launch or two.  The merged kernel saves a launch, which is what a small, launch-latency-bound frame
    // wants (gun1080: 18,670 frames/s merged, 17,440 as two launches).  Its CTAs are twice as heavy, though (64+
    // registers x 256 threads), and beside the previous frame's tile kernel a large frame does better with the two
    // lean launches although they take 4 us longer alone (city4k: 5,690 against 5,590 frames/s).
    // Decided from the last completed frame's counters (a heuristic only -- both orders of filling give the same
    // bins up to order, which no consumer depends on).
    // Larger frames always fill in two launches, the triangles that span many tiles FIRST: they are then at the head of
    // every bin, where the tile kernel's hi-Z wants its occluders.
    const bool mergedFill = prevBig > 0 && prevRefs <= 16384u;
    const unsigned bigCountGrid = (unsigned)(c->smCount * (fullFrame ? c->bigCountCtas : 4));
    // Extra CTAs at the head of the tile kernel's grid, one per part of a split bin beyond the first.  None when bins
    // are not split (every dead CTA still costs a launch slot and two loads: 4,096 of them were 12 % of the city4k
    // grid and half of gun1080's); otherwise what the last frame would have needed plus an eighth, in steps of 256 so
    // that the captured chain keeps its shape.  The scan kernel sizes the parts to fit whatever it is given.
    uint32_t extraCtas = 0u;
    if (splitMin) {
        // parts beyond the first of all split bins <= their entries / part size (published by the last scan kernel)
        const uint32_t need = c->fb[sPrev].hCnt->nSplitEntries / std::max(c->splitPart, 32u);
        extraCtas = std::min<uint32_t>(kMaxExtra, (need + need / 8u + 32u + 255u) / 256u * 256u);
    }
    const bool direct = c->profiling || !c->useGraphs;       // launch by launch: stage events, or graphs switched off

    // Everything of the chain that goes onto the geometry stream, in order.  `live`: issued directly (events may be
    // recorded between the launches); otherwise it is being captured into a graph.
    auto issue_chain = [&](bool live) {
        int n = 0;
        cudaMemsetAsync(B.dCntTiles.p, 0, (64 + 2 * (size_t)nTiles) * sizeof(uint32_t), gs);
        if (live && c->profiling) cudaEventRecord(c->stageEv[0], gs);
        if (fp.nDraws > 0) {
            frustum_draws_kernel<<<(fp.nDraws + 3) / 4, 128, 0, gs>>>(B.dDraws.p, fp.nDraws, U.numerics_fma, B.dCnt, fp.W, fp.H, fp.bandRow0,
                                                                      fp.bandRow1, fp.tilesY, B.dVBlocks.p, B.dTBlocks.p);
            ++n;
            if (!fullFrame) {                              // band context: block tables culled to the band
                band_blocks_kernel<<<fp.nDraws, 128, 0, gs>>>(B.dDraws.p, B.dCnt, B.dVBlocks.p, B.dTBlocks.p);
                ++n;
            }
        }
        if (live) { cudaEventRecord(B.geomStarted, gs); mark(1); }
        if (c->nVBlocks > 0) {
            if (U.numerics_fma)
                vertex_kernel<true><<<gridV, kGeomBlock, 0, gs>>>(B.dDraws.p, B.dVBlocks.p, B.dCnt, B.dClip.p, B.dWn.p);
            else
                vertex_kernel<false><<<gridV, kGeomBlock, 0, gs>>>(B.dDraws.p, B.dVBlocks.p, B.dCnt, B.dClip.p, B.dWn.p);
            ++n;
        }
        if (live) mark(2);
        if (c->nTBlocks > 0) {
            setup_kernel<<<gridT, kGeomBlock, 0, gs>>>(
                B.dDraws.p, B.dTBlocks.p, fp, U, B.dClip.p, B.dWn.p, RecStore{B.dRecsR.p, B.dRecsS.p, (uint32_t)fp.totalT}, B.dClipAttrs.p, B.dRefs.p, B.dBigQ.p,
                B.dTileCount, B.dCnt);
            bin_big_kernel<0><<<bigCountGrid, 256, 0, gs>>>(B.dBigQ.p, B.dCnt, fp, B.dTileCount, B.dTileHasBig);
            n += 2;
        }
        if (live) mark(3);
        if (bandTiles > 0) {
            if (splitMin)
                scan_tiles_kernel<true><<<kScanCtas, 1024, 0, gs>>>(B.dTileCount, B.dTileStart.p, B.dTileCursor.p, B.dOrder.p, fp.bandRow0 * fp.tilesX,
                                                                   bandTiles, B.dCnt, fp.pairCap, splitMin, c->splitPart, (uint32_t*)B.hCnt, B.dTileHasBig, extraCtas);
            else
                scan_tiles_kernel<false><<<kScanCtas, 1024, 0, gs>>>(B.dTileCount, B.dTileStart.p, B.dTileCursor.p, B.dOrder.p, fp.bandRow0 * fp.tilesX,
                                                                    bandTiles, B.dCnt, fp.pairCap, 0u, c->splitPart, (uint32_t*)B.hCnt, B.dTileHasBig, 0u);
            ++n;
        } else {
            publish_counters_kernel<<<1, sizeof(Counters) / 4, 0, gs>>>(B.dCnt, (uint32_t*)B.hCnt);
        }
        if (live) { cudaEventRecord(B.cntReady, gs); mark(4); }    // the counters are final here
        if (fp.totalT > 0 && bandTiles > 0) {
            if (mergedFill) {
                fill_merged_kernel<<<2 * wave, 256, 0, gs>>>(B.dRefs.p, B.dBigQ.p, B.dCnt, fp, B.dTileCursor.p, B.dBins.p, wave);
            } else {
                // grid-stride over the accepted-triangle list (its length is only known on the device)
                bin_big_kernel<1><<<wave, 256, 0, gs>>>(B.dBigQ.p, B.dCnt, fp, B.dTileCursor.p, B.dBins.p);
                fill_bins_kernel<<<fillGrid, 256, 0, gs>>>(B.dRefs.p, B.dCnt, fp, B.dTileCursor.p, B.dBins.p);
                ++n;
            }
            ++n;
        }
        return n;
    };

    if (direct) {
        launches += issue_chain(true);
        B.startedValid = true;
        B.lastKey.clear();
    } else {
        // the shape of this frame's chain on this set
        std::vector<unsigned char> key;
        auto put = [&key](const void* p, size_t n) { key.insert(key.end(), (const unsigned char*)p, (const unsigned char*)p + n); };
        const void* ptrs[] = {B.dDraws.p, B.dVBlocks.p, B.dTBlocks.p, B.dClip.p, B.dWn.p, B.dRecsR.p, B.dRecsS.p, B.dClipAttrs.p,
                              B.dRefs.p, B.dBigQ.p, B.dCntTiles.p, B.dTileStart.p, B.dTileCursor.p, B.dOrder.p, B.dBins.p, B.hCnt};
        const uint32_t ints[] = {gridV, gridT, (uint32_t)c->nVBlocks, (uint32_t)c->nTBlocks, (uint32_t)mergedFill, splitMin, c->splitPart,
                                 bigCountGrid, fillGrid, wave, (uint32_t)nTiles, (uint32_t)bandTiles, extraCtas};
        put(&fp, sizeof(fp)); put(&U, sizeof(U)); put(ptrs, sizeof(ptrs)); put(ints, sizeof(ints));
        bool done = false;
        if (B.graphExec && key == B.graphKey) {
            CK(c, cudaGraphLaunch(B.graphExec, gs));
            launches += B.graphLaunches;
            done = true;
        } else {
            if (B.graphExec) { cudaGraphExecDestroy(B.graphExec); B.graphExec = nullptr; B.graphKey.clear(); }
            if (key == B.lastKey && cudaStreamBeginCapture(gs, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
                const int n = issue_chain(false);
                cudaGraph_t g = nullptr;
                cudaError_t e = cudaStreamEndCapture(gs, &g);
                if (e == cudaSuccess) e = cudaGraphInstantiate(&B.graphExec, g, 0);
                if (g) cudaGraphDestroy(g);
                if (e == cudaSuccess) e = cudaGraphLaunch(B.graphExec, gs);
                if (e == cudaSuccess) {
                    B.graphKey = key;
                    B.graphLaunches = n;
                    launches += n;
                    done = true;
                } else {                                   // never fatal: this context goes back to plain launches
                    cudaGetLastError();
                    if (B.graphExec) { cudaGraphExecDestroy(B.graphExec); B.graphExec = nullptr; }
                    c->useGraphs = false;
                }
            }
        }
        B.lastKey = key;
        if (done) {
            CK(c, cudaEventRecord(B.cntReady, gs));        // graph path: known when the whole chain is done
            B.startedValid = false;                        // (no mid-chain event to order the next frame's chain behind)
        } else {
            launches += issue_chain(true);
            B.startedValid = true;
        }
    }
    mark(5);
    SRB_STAMP(1, gs);
    CK(c, cudaEventRecord(B.geomDone, gs));
    CK(c, cudaStreamWaitEvent(c->stream, B.geomDone, 0));
    SRB_STAMP(2, c->stream);
    RasterArgs A;
    A.draws = B.dDraws.p; A.order = B.dOrder.p; A.wn = B.dWn.p; A.recs.R = B.dRecsR.p; A.recs.S = B.dRecsS.p; A.recs.fan1Base = (uint32_t)fp.totalT;
    A.gkeys = c->dSliceKeys; A.sliceDone = c->dSliceDone; A.bandTiles = bandTiles; A.extraCtas = extraCtas;
    A.clipAttrs = B.dClipAttrs.p; A.tileStart = B.dTileStart.p; A.tileEnd = B.dTileCursor.p;
    A.bins = B.dBins.p; A.color = c->color; A.depth = c->depth; A.cnt = B.dCnt;
    A.colorOut = c->colorTarget ? c->colorTarget : c->color;
    A.color8Out = c->color8Target ? c->color8Target : c->color8;
#ifdef SRB_TILE_TIMING
    if (c->dBinScratch.reserve(((size_t)nTiles + kMaxExtra) * 8) != cudaSuccess) return fail(c, SRB_ERR_CUDA, "dbg alloc");
    CK(c, cudaMemsetAsync(c->dBinScratch.p, 0, ((size_t)nTiles + kMaxExtra) * 32, c->stream));
    A.dbg = (ulonglong4*)c->dBinScratch.p;
#endif
    if (bandTiles > 0) {
        if (c->useVis) {
            mark(6);                       // no sort stage on the visibility path
            const bool out8 = A.color8Out != nullptr;
            const unsigned visGrid = (unsigned)bandTiles + extraCtas;      // the first extraCtas CTAs: parts of split bins
            switch (c->visMode) {
                case 0: if (out8) vis_tile_kernel<0, true><<<visGrid, kVisThreads, 0, c->stream>>>(A, fp, U);
                        else vis_tile_kernel<0, false><<<visGrid, kVisThreads, 0, c->stream>>>(A, fp, U); break;
                case 1: if (out8) vis_tile_kernel<1, true><<<visGrid, kVisThreads, 0, c->stream>>>(A, fp, U);
                        else vis_tile_kernel<1, false><<<visGrid, kVisThreads, 0, c->stream>>>(A, fp, U); break;
                case 2: if (out8) vis_tile_kernel<2, true><<<visGrid, kVisThreads, 0, c->stream>>>(A, fp, U);
                        else vis_tile_kernel<2, false><<<visGrid, kVisThreads, 0, c->stream>>>(A, fp, U); break;
                default: if (out8) vis_tile_kernel<3, true><<<visGrid, kVisThreads, 0, c->stream>>>(A, fp, U);
                         else vis_tile_kernel<3, false><<<visGrid, kVisThreads, 0, c->stream>>>(A, fp, U); break;
            }
            ++launches;
        } else {
            sort_bins_kernel<<<bandTiles, 256, 0, c->stream>>>(B.dTileStart.p, B.dTileCursor.p, B.dBins.p,
                                                              c->dBinScratch.p, fp.bandRow0 * fp.tilesX, B.dCnt);
            mark(6);
            ordered_tile_kernel<<<bandTiles, 256, 0, c->stream>>>(A, fp, U);
            launches += 2;
        }
    }
    mark(7);
    SRB_STAMP(3, c->stream);
#ifdef SRB_TIMELINE
    ++c->timelineFrame;
#endif
    CK(c, cudaGetLastError());
    CK(c, cudaEventRecord(B.rasterDone, c->stream));
    B.rasterRecorded = true;
    CK(c, cudaEventRecord(c->ev1, c->stream));
    c->stats.kernel_launches = launches;
    c->stats.pipeline_used = c->useVis ? SRB_PIPELINE_VISIBILITY : SRB_PIPELINE_ORDERED;
    c->framePending = true;
    return SRB_OK;
}

// Settle the pending frame: wait until its geometry stages are done (the counters are final after the scan kernel),
// and if a device-side capacity overflowed, grow the buffers and run the frame again -- every raster kernel of the
// overflowed attempt returns at once, so the framebuffer is untouched, also for frames that do not clear.
// waitRaster: additionally wait for the frame's pixels (read-backs, accessors, statistics with timings).
int complete_frame(SrbContext* c, bool waitRaster = true) {
    if (!c->framePending) {
        if (waitRaster) CK(c, cudaStreamSynchronize(c->stream));
        return SRB_OK;
    }
    CK(c, cudaEventSynchronize(c->fb[c->slot].cntReady));
    for (int attempt = 0; attempt < 8; ++attempt) {
        const Counters& k = *c->fb[c->slot].hCnt;
        if (!k.overflow) break;
        // (cudaFree inside reserve() waits for the device: nothing is freed under a running kernel.)
        for (int i = 0; i < c->nSets; ++i) {
            SrbContext::FrameBufs& b = c->fb[i];
            if (k.overflow & 1u) CK(c, b.dBins.reserve((size_t)k.nPairs + k.nPairs / 8 + 1024));
            if (k.overflow & 2u) CK(c, b.dClipAttrs.reserve(std::max<size_t>(4 * (size_t)k.nClipped, 4096)));
            if (k.overflow & 4u) CK(c, b.dBigQ.reserve((size_t)k.nBig + k.nBig / 4 + 1024));
        }
        if (k.overflow & 1u) CK(c, c->dBinScratch.reserve(c->fb[0].dBins.cap));
        refresh_caps(c);
        int rc = launch_frame(c);
        if (rc != SRB_OK) return rc;
        CK(c, cudaEventSynchronize(c->fb[c->slot].cntReady));
    }
    if (c->fb[c->slot].hCnt->overflow) return fail(c, SRB_ERR_CUDA, "bin capacity overflow persisted after regrow");
    const Counters& k = *c->fb[c->slot].hCnt;
    c->stats.draws_visible = k.nVisibleDraws;
    c->stats.triangles_setup = k.nRefs;
    c->stats.split_parts = (int32_t)std::min<uint32_t>(k.nExtra, 0x7fffffffu);
    c->stats.triangles_no_coverage = (int32_t)k.nNoCover;
    c->stats.tile_pairs = k.nPairs;
    c->stats.clipped_triangles = k.nClippedTris;
    if (!waitRaster) return SRB_OK;
    CK(c, cudaStreamSynchronize(c->stream));
    c->framePending = false;
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) c->stats.gpu_ms = ms;
    if (c->profiling)
        for (int i = 0; i < SRB_N_STAGES; ++i) {
            float t = 0.0f;
            c->stageMs[i] = cudaEventElapsedTime(&t, c->stageEv[i], c->stageEv[i + 1]) == cudaSuccess ? t : 0.0f;
        }
    return SRB_OK;
}

}  // namespace

// ===========================================================================
// C ABI
// ===========================================================================
extern "C" {

int srb_version(void) { return 100; }

const char* srb_last_error(SrbContext* ctx) { return ctx ? ctx->err.c_str() : g_createError.c_str(); }

int srb_create(int device, int width, int height, SrbContext** out_ctx) {
    if (!out_ctx) return fail(nullptr, SRB_ERR_ARG, "out_ctx is null");
    *out_ctx = nullptr;
    if (width <= 0 || height <= 0)      // Rasterizer.InitializeTileLocks throws ArgumentException (:71-74)
        return fail(nullptr, SRB_ERR_ARG, "Width and height must be positive non-zero values.");
    if (width > 65535 || height > 65535) return fail(nullptr, SRB_ERR_ARG, "render target larger than 65535");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0)
        return fail(nullptr, SRB_ERR_CUDA, std::string("no CUDA device (srb200 has no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(nullptr, SRB_ERR_ARG, "bad device index");
    SrbContext* c = new SrbContext();
    c->device = device;
    std::memset(&c->stats, 0, sizeof(c->stats));
    std::memset(&c->fp, 0, sizeof(c->fp));
    default_uniforms(c->uniforms);
    auto bail = [&](const char* what, cudaError_t err) {
        g_createError = std::string(what) + ": " + cudaGetErrorString(err);
        delete c;
        return SRB_ERR_CUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
    cudaDeviceGetAttribute(&c->smCount, cudaDevAttrMultiProcessorCount, device);
    if (c->smCount <= 0) c->smCount = 148;
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
    if ((e = cudaEventCreate(&c->ev0)) != cudaSuccess) return bail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&c->ev1)) != cudaSuccess) return bail("cudaEventCreate", e);
    int prioLo = 0, prioHi = 0;       // geometry kernels are short and latency-bound: let their CTAs go ahead of the queued
    cudaDeviceGetStreamPriorityRange(&prioLo, &prioHi);     // tile CTAs of the previous frame
    for (int s = 0; s < 2; ++s)
        if ((e = cudaEventCreateWithFlags(&c->stagingDone[s], cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
    for (int s = 0; s < SrbContext::kMaxSets; ++s) {
        if ((e = cudaStreamCreateWithPriority(&c->fb[s].geomStream, cudaStreamNonBlocking, prioHi)) != cudaSuccess) return bail("cudaStreamCreate", e);
        if ((e = cudaEventCreateWithFlags(&c->fb[s].geomDone, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
        if ((e = cudaEventCreateWithFlags(&c->fb[s].geomStarted, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
        if ((e = cudaEventCreateWithFlags(&c->fb[s].rasterDone, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
        if ((e = cudaEventCreateWithFlags(&c->fb[s].cntReady, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);

        if ((e = cudaMallocHost((void**)&c->fb[s].hCnt, sizeof(Counters))) != cudaSuccess) return bail("cudaMallocHost", e);
        std::memset(c->fb[s].hCnt, 0, sizeof(Counters));
    }
    if ((e = cudaMalloc((void**)&c->dSliceKeys, (size_t)kMaxExtra * kTile * kTile * sizeof(unsigned long long))) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMalloc((void**)&c->dSliceDone, (size_t)kMaxExtra * sizeof(uint32_t))) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMemsetAsync(c->dSliceKeys, 0, (size_t)kMaxExtra * kTile * kTile * sizeof(unsigned long long), c->stream)) != cudaSuccess) return bail("cudaMemset", e);
    if ((e = cudaMemsetAsync(c->dSliceDone, 0, (size_t)kMaxExtra * sizeof(uint32_t), c->stream)) != cudaSuccess) return bail("cudaMemset", e);
#ifdef SRB_TIMELINE
    if ((e = cudaMalloc((void**)&c->dTimeline, (size_t)kTimelineFrames * 4 * sizeof(unsigned long long))) != cudaSuccess) return bail("cudaMalloc", e);
#endif
    int rc = alloc_framebuffer(c, width, height);
    if (rc != SRB_OK) { g_createError = c->err; delete c; return rc; }
    *out_ctx = c;
    return SRB_OK;
}

int srb_destroy(SrbContext* c) {
    if (!c) return SRB_ERR_ARG;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    close_color_target(c);
    close_color8_target(c);
    barrier_close(c);
    if (c->barStream) { cudaStreamSynchronize(c->barStream); cudaStreamDestroy(c->barStream); }
    if (c->barFrame) cudaEventDestroy(c->barFrame);
    for (auto& e : c->barDone) if (e) cudaEventDestroy(e);
    if (c->barRow) cudaFree(c->barRow);
    if (c->barTimedOut) cudaFreeHost(c->barTimedOut);
    if (c->color8) cudaFree(c->color8);
    for (auto& m : c->meshes) if (m.alive) { cudaFree(m.vtx); cudaFree(m.idx); cudaFree(m.blockBoxes); cudaFree(m.blockVRange); }
    for (auto& t : c->textures) if (t.alive) cudaFree(t.px);

    if (c->dSliceKeys) cudaFree(c->dSliceKeys);
    if (c->dSliceDone) cudaFree(c->dSliceDone);
    c->dRayScratch.release();
    c->dBinScratch.release();
    for (auto& b : c->fb) {
        b.dDraws.release(); b.dWn.release(); b.dRecsR.release(); b.dRecsS.release(); b.dClipAttrs.release(); b.dTileStart.release();
        b.dTileCursor.release(); b.dBins.release(); b.dOrder.release();
        b.dVBlocks.release(); b.dTBlocks.release(); b.dClip.release(); b.dRefs.release(); b.dBigQ.release(); b.dCntTiles.release();
        if (b.geomStream) { cudaStreamSynchronize(b.geomStream); cudaStreamDestroy(b.geomStream); }
        if (b.geomDone) cudaEventDestroy(b.geomDone);
        if (b.geomStarted) cudaEventDestroy(b.geomStarted);
        if (b.graphExec) cudaGraphExecDestroy(b.graphExec);
        if (b.rasterDone) cudaEventDestroy(b.rasterDone);
        if (b.cntReady) cudaEventDestroy(b.cntReady);
        if (b.hCnt) cudaFreeHost(b.hCnt);
    }
    for (int s = 0; s < 2; ++s) {
        if (c->stagingDone[s]) cudaEventDestroy(c->stagingDone[s]);
        if (c->hDraws[s]) cudaFreeHost(c->hDraws[s]);
    }
    if (c->copyStream) { cudaStreamSynchronize(c->copyStream); cudaStreamDestroy(c->copyStream); }
    for (int s = 0; s < 2; ++s) {
        c->dPresent[s].release();
        if (c->presentReady[s]) cudaEventDestroy(c->presentReady[s]);
        if (c->presentCopied[s]) cudaEventDestroy(c->presentCopied[s]);
    }
    if (c->color) cudaFree(c->color);
    if (c->depth) cudaFree(c->depth);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    for (auto& e : c->stageEv) if (e) cudaEventDestroy(e);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return SRB_OK;
}

int srb_resize(SrbContext* c, int width, int height) {
    if (!c) return SRB_ERR_ARG;
    if (width <= 0 || height <= 0) return fail(c, SRB_ERR_ARG, "Width and height must be positive non-zero values.");
    if (width > 65535 || height > 65535) return fail(c, SRB_ERR_ARG, "render target larger than 65535");
    if (c->inFrame) return fail(c, SRB_ERR_STATE, "resize inside a frame");
    CK(c, cudaSetDevice(c->device));
    int rc = complete_frame(c);
    if (rc != SRB_OK) return rc;
    c->bandSet = false;
    return alloc_framebuffer(c, width, height);
}

int srb_get_size(SrbContext* c, int* w, int* h) {
    if (!c) return SRB_ERR_ARG;
    if (w) *w = c->W;
    if (h) *h = c->H;
    return SRB_OK;
}

int srb_set_band(SrbContext* c, int r0, int r1) {
    if (!c) return SRB_ERR_ARG;
    if (r0 < 0 || r1 < r0 || r1 > c->tilesY) return fail(c, SRB_ERR_ARG, "band rows out of range");
    c->band0 = r0; c->band1 = r1; c->bandSet = true;
    return SRB_OK;
}

int srb_export_color_handle(SrbContext* c, uint8_t handle_out[SRB_IPC_HANDLE_BYTES]) {
    if (!c || !handle_out) return SRB_ERR_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == SRB_IPC_HANDLE_BYTES, "IPC handle size");
    CK(c, cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    CK(c, cudaIpcGetMemHandle(&h, c->color));
    std::memcpy(handle_out, &h, sizeof(h));
    return SRB_OK;
}

int srb_set_color_target(SrbContext* c, const uint8_t handle[SRB_IPC_HANDLE_BYTES]) {
    if (!c) return SRB_ERR_ARG;
    if (c->inFrame) return fail(c, SRB_ERR_STATE, "colour target change inside a frame");
    CK(c, cudaSetDevice(c->device));
    int rc = complete_frame(c);
    if (rc != SRB_OK) return rc;
    close_color_target(c);
    if (!handle) return SRB_OK;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    CK(c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->colorTarget = (float4*)p;
    return SRB_OK;
}

/* RGBA8 band gather (payload float4 or RGBA8 selectable). */
int srb_gather_rgba8_export(SrbContext* c, uint8_t handle_out[SRB_IPC_HANDLE_BYTES]) {
    if (!c || !handle_out) return SRB_ERR_ARG;
    if (c->inFrame) return fail(c, SRB_ERR_STATE, "gather frame change inside a frame");
    CK(c, cudaSetDevice(c->device));
    int rc = complete_frame(c);
    if (rc != SRB_OK) return rc;
    if (!c->color8) {
        CK(c, cudaMalloc((void**)&c->color8, (size_t)c->W * c->H * sizeof(uint32_t)));
        CK(c, cudaMemset(c->color8, 0, (size_t)c->W * c->H * sizeof(uint32_t)));
    }
    cudaIpcMemHandle_t h;
    CK(c, cudaIpcGetMemHandle(&h, c->color8));
    std::memcpy(handle_out, &h, sizeof(h));
    return SRB_OK;
}

int srb_gather_rgba8_target(SrbContext* c, const uint8_t handle[SRB_IPC_HANDLE_BYTES]) {
    if (!c) return SRB_ERR_ARG;
    if (c->inFrame) return fail(c, SRB_ERR_STATE, "gather target change inside a frame");
    CK(c, cudaSetDevice(c->device));
    int rc = complete_frame(c);
    if (rc != SRB_OK) return rc;
    close_color8_target(c);
    if (!handle) return SRB_OK;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    CK(c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->color8Target = (uint32_t*)p;
    return SRB_OK;
}

int srb_read_gathered_rgba8(SrbContext* c, uint8_t* out) {
    if (!c || !out) return SRB_ERR_ARG;
    if (!c->color8) return fail(c, SRB_ERR_STATE, "no RGBA8 gather frame on this context (srb_gather_rgba8_export first)");
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    CK(c, cudaMemcpy(out, c->color8, (size_t)c->W * c->H * 4, cudaMemcpyDeviceToHost));
    return SRB_OK;
}

void* srb_device_gathered_rgba8_ptr(SrbContext* c) { return c ? (void*)c->color8 : nullptr; }

int srb_get_tile_row_pairs(SrbContext* c, uint32_t* out, int n) {
    if (!c || !out || n < 0) return SRB_ERR_ARG;
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    if (!c->haveFrame) return fail(c, SRB_ERR_STATE, "no completed frame");
    const int nTiles = c->tilesX * c->tilesY;
    std::vector<uint32_t> h((size_t)nTiles);
    CK(c, cudaMemcpy(h.data(), c->fb[c->slot].dTileCount, (size_t)nTiles * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    for (int r = 0; r < n && r < c->tilesY; ++r) {
        uint64_t s = 0;
        for (int x = 0; x < c->tilesX; ++x) s += h[(size_t)r * c->tilesX + x];
        out[r] = (uint32_t)std::min<uint64_t>(s, 0xffffffffu);
    }
    return SRB_OK;
}

int srb_set_bin_split(SrbContext* c, int min_entries_full_frame, int min_entries_band, int part_entries) {
    if (!c) return SRB_ERR_ARG;
    if (min_entries_full_frame < 0 || min_entries_band < 0 || part_entries < 32 || part_entries % 32 != 0 ||
        (min_entries_full_frame != 0 && min_entries_full_frame < 64) || (min_entries_band != 0 && min_entries_band < 64))
        return fail(c, SRB_ERR_ARG, "bin split: thresholds are 0 (off) or >= 64 entries, parts a multiple of 32 entries");
    c->splitMinFull = (uint32_t)min_entries_full_frame;
    c->splitMinBand = (uint32_t)min_entries_band;
    c->splitPart = (uint32_t)part_entries;
    return SRB_OK;
}

int srb_set_graphs(SrbContext* c, int enabled) {
    if (!c) return SRB_ERR_ARG;
    c->useGraphs = enabled != 0;
    return SRB_OK;
}

int srb_set_pipeline_depth(SrbContext* c, int sets) {
    if (!c) return SRB_ERR_ARG;
    if (sets != 0 && (sets < 2 || sets > SrbContext::kMaxSets)) return fail(c, SRB_ERR_ARG, "pipeline depth must be 0 (automatic) or 2..4 buffer sets");
    c->pipeDepth = sets;
    return SRB_OK;
}

int srb_set_pipeline(SrbContext* c, int p) {
    if (!c) return SRB_ERR_ARG;
    if (p < 0 || p > 2) return fail(c, SRB_ERR_ARG, "bad pipeline");
    c->pipeline = p;
    return SRB_OK;
}

int srb_texture_create(SrbContext* c, const uint8_t* rgba8, int w, int h, SrbTexture* out) {
    if (!c || !out) return SRB_ERR_ARG;
    if (!rgba8 || w <= 0 || h <= 0) return fail(c, SRB_ERR_ARG, "bad texture");
    CK(c, cudaSetDevice(c->device));
    TexRes t;
    t.w = w; t.h = h;
    size_t n = (size_t)w * h;
    t.opaque = true;
    for (size_t i = 0; i < n; ++i) if (rgba8[4 * i + 3] != 255) { t.opaque = false; break; }
    CK(c, cudaMalloc((void**)&t.px, n * 4));
    CK(c, cudaMemcpy(t.px, rgba8, n * 4, cudaMemcpyHostToDevice));
    t.alive = true;
    std::lock_guard<std::mutex> lk(c->mu);
    c->textures.push_back(t);
    *out = (SrbTexture)c->textures.size();
    return SRB_OK;
}

int srb_texture_destroy(SrbContext* c, SrbTexture tex) {
    if (!c) return SRB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    if (tex <= 0 || (size_t)tex > c->textures.size() || !c->textures[tex - 1].alive)
        return fail(c, SRB_ERR_HANDLE, "unknown texture handle");
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaFree(c->textures[tex - 1].px);
    c->textures[tex - 1] = TexRes();
    c->haveFrame = false;
    return SRB_OK;
}

int srb_texture_sample(SrbContext* c, SrbTexture tex, const float* uv, int n, int filter, float* out) {
    if (!c || !uv || !out || n < 0) return SRB_ERR_ARG;
    if (n == 0) return SRB_OK;
    TexRes t;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        if (tex <= 0 || (size_t)tex > c->textures.size() || !c->textures[tex - 1].alive)
            return fail(c, SRB_ERR_HANDLE, "unknown texture handle");
        t = c->textures[tex - 1];
    }
    CK(c, cudaSetDevice(c->device));
    float* d = nullptr;
    CK(c, cudaMalloc((void**)&d, (size_t)n * 24));
    float4* dOut = (float4*)d;
    float2* dUv = (float2*)(d + (size_t)n * 4);
    cudaError_t e = cudaMemcpyAsync(dUv, uv, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        texture_sample_kernel<<<(n + 127) / 128, 128, 0, c->stream>>>(t.px, t.w, t.h, dUv, n, filter, dOut);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dOut, (size_t)n * 16, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if (e != cudaSuccess) return fail(c, SRB_ERR_CUDA, std::string("srb_texture_sample: ") + cudaGetErrorString(e));
    return SRB_OK;
}

int srb_mesh_create(SrbContext* c, const float* vertices, int nV, const uint16_t* indices, int nI, SrbMesh* out) {
    if (!c || !out) return SRB_ERR_ARG;
    if (!vertices || !indices || nV <= 0 || nI < 0) return fail(c, SRB_ERR_ARG, "bad mesh arrays");
    if (nV > 65536) return fail(c, SRB_ERR_ARG, "mesh exceeds ushort index range (ModelLoader.cs:47)");
    for (int i = 0; i < nI; ++i)
        if (indices[i] >= nV) return fail(c, SRB_ERR_ARG, "index out of range (IndexOutOfRangeException in the reference)");
    CK(c, cudaSetDevice(c->device));
    MeshRes m;
    m.nV = nV; m.nI = nI;
    m.opaque = true;
    for (int i = 0; i < nV; ++i) if (vertices[12 * i + 11] != 1.0f) { m.opaque = false; break; }
    CK(c, cudaMalloc((void**)&m.vtx, (size_t)nV * 48));
    CK(c, cudaMalloc((void**)&m.idx, std::max<size_t>((size_t)nI * 2, 16)));
    CK(c, cudaMemcpy(m.vtx, vertices, (size_t)nV * 48, cudaMemcpyHostToDevice));
    if (nI > 0) CK(c, cudaMemcpy(m.idx, indices, (size_t)nI * 2, cudaMemcpyHostToDevice));
    {   // object-space bounding box (centre, half extents) of every run of kGeomBlock triangles: block culling of band
        // contexts.  (A bounding sphere kept 82 % of city4k's blocks on the band that owns the upper half of the
        // screen, the box keeps 39 %, within a point of the blocks' true screen extent: walls and floors are flat.)
        const int nT = nI / 3, nB = (nT + kGeomBlock - 1) / kGeomBlock;
        std::vector<float4> box(2 * (size_t)std::max(nB, 1), make_float4(0.0f, 0.0f, 0.0f, 0.0f));
        for (int b = 0; b < nB; ++b) {
            const int t0 = b * kGeomBlock, t1 = std::min(nT, t0 + kGeomBlock);
            float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
            bool finite = true;
            for (int i = 3 * t0; i < 3 * t1; ++i)
                for (int k = 0; k < 3; ++k) {
                    const float v = vertices[12 * (size_t)indices[i] + k];
                    finite = finite && std::isfinite(v);
                    lo[k] = std::min(lo[k], v); hi[k] = std::max(hi[k], v);
                }
            if (!finite) { box[2 * b + 1] = make_float4(3.0e38f, 3.0e38f, 3.0e38f, 0.0f); continue; }   // never culled
            float ctr[3], half[3];
            for (int k = 0; k < 3; ++k) {
                const double cd = 0.5 * ((double)lo[k] + hi[k]);
                ctr[k] = (float)cd;
                // half extent about the ROUNDED centre, rounded up: the box contains every vertex
                half[k] = (float)((std::max((double)hi[k] - ctr[k], (double)ctr[k] - lo[k])) * 1.0001 + 1e-6);
            }
            box[2 * b] = make_float4(ctr[0], ctr[1], ctr[2], 0.0f);
            box[2 * b + 1] = make_float4(half[0], half[1], half[2], 0.0f);
        }
        CK(c, cudaMalloc((void**)&m.blockBoxes, box.size() * sizeof(float4)));
        CK(c, cudaMemcpy(m.blockBoxes, box.data(), box.size() * sizeof(float4), cudaMemcpyHostToDevice));
        // lowest / highest vertex index of every run (band contexts transform only the vertex runs a kept block uses)
        std::vector<int2> vr((size_t)std::max(nB, 1), make_int2(0, 0));
        for (int b = 0; b < nB; ++b) {
            const int t0 = b * kGeomBlock, t1 = std::min(nT, t0 + kGeomBlock);
            int lo = 65535, hi = 0;
            for (int i = 3 * t0; i < 3 * t1; ++i) { lo = std::min<int>(lo, indices[i]); hi = std::max<int>(hi, indices[i]); }
            vr[b] = make_int2(std::min(lo, hi), hi);
        }
        CK(c, cudaMalloc((void**)&m.blockVRange, vr.size() * sizeof(int2)));
        CK(c, cudaMemcpy(m.blockVRange, vr.data(), vr.size() * sizeof(int2), cudaMemcpyHostToDevice));
    }
    m.alive = true;
    std::lock_guard<std::mutex> lk(c->mu);
    c->meshes.push_back(m);
    *out = (SrbMesh)c->meshes.size();
    return SRB_OK;
}

int srb_mesh_destroy(SrbContext* c, SrbMesh mesh) {
    if (!c) return SRB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    if (mesh <= 0 || (size_t)mesh > c->meshes.size() || !c->meshes[mesh - 1].alive)
        return fail(c, SRB_ERR_HANDLE, "unknown mesh handle");
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaFree(c->meshes[mesh - 1].vtx);
    cudaFree(c->meshes[mesh - 1].idx);
    cudaFree(c->meshes[mesh - 1].blockBoxes);
    cudaFree(c->meshes[mesh - 1].blockVRange);
    c->meshes[mesh - 1] = MeshRes();
    c->haveFrame = false;
    return SRB_OK;
}

int srb_set_uniforms(SrbContext* c, const SrbUniforms* u) {
    if (!c || !u) return SRB_ERR_ARG;
    if (u->n_extra_lights < 0 || u->n_extra_lights > 4) return fail(c, SRB_ERR_ARG, "n_extra_lights must be 0..4");
    if (u->tex_filter < 0 || u->tex_filter > 1) return fail(c, SRB_ERR_ARG, "tex_filter must be 0 or 1");
    if (u->debug_mode != SRB_DEBUG_NONE && u->debug_mode != SRB_DEBUG_WIREFRAME)
        return fail(c, SRB_ERR_ARG, "debug_mode must be 0 (None) or 1 (Wireframe)");
    c->uniforms = *u;
    return SRB_OK;
}

int srb_begin_frame(SrbContext* c, const float clear_rgba[4], uint32_t clear_flags) {
    if (!c) return SRB_ERR_ARG;
    if (c->inFrame) return fail(c, SRB_ERR_STATE, "begin_frame inside a frame");
    if ((clear_flags & SRB_CLEAR_COLOR) && !clear_rgba) return fail(c, SRB_ERR_ARG, "clear colour missing");
    if (clear_rgba) std::memcpy(c->clear, clear_rgba, sizeof(c->clear));
    c->clearFlags = clear_flags;
    c->queue.clear();
    c->callCounter = 0;
    c->inFrame = true;
    return SRB_OK;
}

static int queue_draw(SrbContext* c, const SrbDrawCmd& cmd) {
    if (cmd.mesh <= 0 || (size_t)cmd.mesh > c->meshes.size() || !c->meshes[cmd.mesh - 1].alive)
        return fail(c, SRB_ERR_HANDLE, "unknown mesh handle");
    if (cmd.texture < 0 || (size_t)cmd.texture > c->textures.size() ||
        (cmd.texture > 0 && !c->textures[cmd.texture - 1].alive))
        return fail(c, SRB_ERR_HANDLE, "unknown texture handle");
    if (cmd.cull < 0 || cmd.cull > 2 || cmd.depth_test < 0 || cmd.depth_test > 7 || cmd.blend < 0 || cmd.blend > 3)
        return fail(c, SRB_ERR_ARG, "bad cull/depth/blend enum");
    QueuedDraw q;
    q.cmd = cmd;
    q.order = cmd.submission_index >= 0 ? (int64_t)cmd.submission_index : (int64_t)0x40000000 + c->callCounter;
    ++c->callCounter;
    c->queue.push_back(q);
    return SRB_OK;
}

int srb_draw(SrbContext* c, SrbMesh mesh, const float model[16], const float view[16], const float proj[16],
             SrbTexture tex, int cull, int depth_test, int blend, int submission_index) {
    if (!c || !model || !view || !proj) return SRB_ERR_ARG;
    SrbDrawCmd cmd;
    std::memset(&cmd, 0, sizeof(cmd));
    cmd.mesh = mesh; cmd.texture = tex;
    std::memcpy(cmd.model, model, 64); std::memcpy(cmd.view, view, 64); std::memcpy(cmd.proj, proj, 64);
    cmd.cull = cull; cmd.depth_test = depth_test; cmd.blend = blend;
    cmd.submission_index = submission_index;
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->inFrame) return fail(c, SRB_ERR_STATE, "draw outside begin/end frame");
    return queue_draw(c, cmd);
}

int srb_submit_draws(SrbContext* c, const SrbDrawCmd* cmds, int n) {
    if (!c || (!cmds && n > 0) || n < 0) return SRB_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->inFrame) return fail(c, SRB_ERR_STATE, "draw outside begin/end frame");
    c->queue.reserve(c->queue.size() + n);
    for (int i = 0; i < n; ++i) {
        int rc = queue_draw(c, cmds[i]);
        if (rc != SRB_OK) return rc;
    }
    return SRB_OK;
}

int srb_end_frame(SrbContext* c) {
    if (!c) return SRB_ERR_ARG;
    if (!c->inFrame) return fail(c, SRB_ERR_STATE, "end_frame without begin_frame");
    c->inFrame = false;
    CK(c, cudaSetDevice(c->device));
    // The previous frame only has to be past its geometry stages (capacities settled, or grown and re-run): its
    // tile kernel keeps running while this frame is prepared.  Everything queued below is ordered behind it on
    // c->stream, and the pinned staging buffers alternate.
    int rc = complete_frame(c, false);
    if (rc != SRB_OK) return rc;
    std::stable_sort(c->queue.begin(), c->queue.end(),
                     [](const QueuedDraw& a, const QueuedDraw& b) { return a.order < b.order; });
    const int nDraws = (int)c->queue.size();
    const int hs = c->stagingSlot;
    c->stagingSlot ^= 1;
    if (c->stagingBusy[hs]) {                    // an upload of the frame before last
        CK(c, cudaEventSynchronize(c->stagingDone[hs]));
        c->stagingBusy[hs] = false;
    }
    if ((size_t)nDraws + 1 > c->hDrawsCap[hs]) {
        if (c->hDraws[hs]) cudaFreeHost(c->hDraws[hs]);
        c->hDraws[hs] = nullptr;
        size_t cap = (size_t)nDraws + nDraws / 2 + 64;
        CK(c, cudaMallocHost((void**)&c->hDraws[hs], cap * sizeof(DrawRec)));
        c->hDrawsCap[hs] = cap;
    }
    DrawRec* const hDraws = c->hDraws[hs];
    size_t nVB = 0, nTB = 0;
    for (int i = 0; i < nDraws; ++i) {
        const MeshRes& m = c->meshes[c->queue[i].cmd.mesh - 1];
        nVB += ((size_t)m.nV + kGeomBlock - 1) / kGeomBlock;
        nTB += ((size_t)(m.nI / 3) + kGeomBlock - 1) / kGeomBlock;
    }
    int64_t totalV = 0, totalT = 0;
    bool eligible = (c->clearFlags & SRB_CLEAR_COLOR) && (c->clearFlags & SRB_CLEAR_DEPTH);
    int commonDepth = -1;
    for (int i = 0; i < nDraws; ++i) {
        const SrbDrawCmd& q = c->queue[i].cmd;
        const MeshRes& m = c->meshes[q.mesh - 1];
        DrawRec& d = hDraws[i];
        std::memcpy(d.model, q.model, 64); std::memcpy(d.view, q.view, 64); std::memcpy(d.proj, q.proj, 64);
        std::memcpy(d.sphere, q.sphere, 16);
        d.vtx = m.vtx; d.idx = m.idx;
        d.tex = nullptr; d.tex_w = d.tex_h = 0;
        bool texOpaque = true;
        if (q.texture > 0) {
            const TexRes& t = c->textures[q.texture - 1];
            d.tex = t.px; d.tex_w = t.w; d.tex_h = t.h;
            texOpaque = t.opaque;
        }
        d.n_vertices = m.nV; d.n_triangles = m.nI / 3;
        d.vbase = (int32_t)totalV; d.tbase = (int32_t)totalT;
        d.cull = q.cull; d.depth_test = q.depth_test; d.blend = q.blend;
        d.flags = q.flags; d.visible = 1; d._pad = 0;
        d.tblockBoxes = m.blockBoxes;
        d.tblockVRange = m.blockVRange;
        for (int k = 0; k < 4; ++k) d.planes[k] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
        totalV += m.nV; totalT += m.nI / 3;
        // Visibility path eligibility: extreme-depth winner with
        // alpha provably 1, one depth function for the whole frame.
        if (q.depth_test < SRB_DEPTH_LESS || q.depth_test > SRB_DEPTH_GREATER_EQUAL) eligible = false;
        if (commonDepth < 0) commonDepth = q.depth_test; else if (commonDepth != q.depth_test) eligible = false;
        if (q.blend != SRB_BLEND_ALPHA && q.blend != SRB_BLEND_NONE) eligible = false;
        if (!m.opaque || !texOpaque) eligible = false;
    }
    if (totalV > 0x3fffffff || totalT > 0x3fffffff) return fail(c, SRB_ERR_ARG, "frame too large (>2^30 vertices or triangles)");
    if (c->colorTarget && !(c->clearFlags & SRB_CLEAR_COLOR))
        return fail(c, SRB_ERR_STATE, "a remote colour target needs frames that clear colour");
    if (c->uniforms.debug_mode == SRB_DEBUG_WIREFRAME) eligible = false;   // DrawLine: depth 1/d, W != 0 writes -- ordered replay only
    if (c->pipeline == SRB_PIPELINE_ORDERED) eligible = false;
    if (c->pipeline == SRB_PIPELINE_VISIBILITY && !eligible && nDraws > 0)
        return fail(c, SRB_ERR_STATE, "frame is not eligible for the visibility pipeline");
    c->useVis = eligible || nDraws == 0;
    if (!((c->clearFlags & SRB_CLEAR_COLOR) && (c->clearFlags & SRB_CLEAR_DEPTH))) c->useVis = false;
    switch (commonDepth) {
        case SRB_DEPTH_LESS: c->visMode = 1; break;
        case SRB_DEPTH_GREATER_EQUAL: c->visMode = 2; break;
        case SRB_DEPTH_GREATER: c->visMode = 3; break;
        default: c->visMode = 0; break;
    }
    const int nTiles = c->tilesX * c->tilesY;
    for (int i = 0; i < c->nSets; ++i) {
        SrbContext::FrameBufs& b = c->fb[i];
        CK(c, b.dDraws.reserve((size_t)nDraws + 1));
        CK(c, b.dVBlocks.reserve(nVB + 1));
        CK(c, b.dTBlocks.reserve(nTB + 1));
        CK(c, b.dOrder.reserve((size_t)nTiles + kMaxExtra));
        CK(c, b.dClip.reserve((size_t)totalV + 1));
        CK(c, b.dWn.reserve((size_t)totalV + 1));
        CK(c, b.dRecsR.reserve(3 * (2 * (size_t)totalT + 2)));
        CK(c, b.dRecsS.reserve(2 * (2 * (size_t)totalT + 2)));
        CK(c, b.dClipAttrs.reserve(4096));
        CK(c, b.dTileStart.reserve(nTiles));
        CK(c, b.dTileCursor.reserve(nTiles));
        CK(c, b.dBins.reserve(std::max<size_t>(1u << 20, 3 * (size_t)totalT)));
        CK(c, b.dRefs.reserve(2 * (size_t)totalT + 64));
        CK(c, b.dBigQ.reserve(1u << 16));
        CK(c, b.dCntTiles.reserve(64 + 2 * (size_t)nTiles));
        b.dCnt = reinterpret_cast<Counters*>(b.dCntTiles.p);
        b.dTileCount = b.dCntTiles.p + 64;
        b.dTileHasBig = b.dTileCount + nTiles;
    }
    CK(c, c->dBinScratch.reserve(c->fb[0].dBins.cap));
    c->nVBlocks = (int)nVB; c->nTBlocks = (int)nTB;
    FrameParams& fp = c->fp;
    fp.W = c->W; fp.H = c->H; fp.tilesX = c->tilesX; fp.tilesY = c->tilesY;
    fp.bandRow0 = c->band0; fp.bandRow1 = c->band1;
    fp.nDraws = nDraws; fp.totalV = (int)totalV; fp.totalT = (int)totalT;
    refresh_caps(c);
    std::memcpy(fp.clearColor, c->clear, 16);
    fp.loadColor = (c->clearFlags & SRB_CLEAR_COLOR) ? 0 : 1;
    fp.loadDepth = (c->clearFlags & SRB_CLEAR_DEPTH) ? 0 : 1;
    fp.depthMode = commonDepth;
    fp.wireframe = c->uniforms.debug_mode == SRB_DEBUG_WIREFRAME ? 1 : 0;
    c->stats.draws_submitted = nDraws;
    c->stats.triangles_submitted = totalT;
    c->haveFrame = true;
    c->tableSlot = hs;                // launch_frame uploads the table to whichever set renders next (and to the others
    ++c->frameSerial;                 // when they are first used for this frame, e.g. by srb_replay_frame)
    return launch_frame(c);
}

int srb_replay_frame(SrbContext* c) {
    if (!c) return SRB_ERR_ARG;
    if (!c->haveFrame || c->inFrame) return fail(c, SRB_ERR_STATE, "no resident frame to replay");
    if (c->fp.loadColor || c->fp.loadDepth) return fail(c, SRB_ERR_STATE, "cannot replay a frame that preserves the framebuffer");
    CK(c, cudaSetDevice(c->device));
    if (c->framePending && c->fb[c->slot].hCnt->overflow) {       // conservative: settle capacities first
        int rc = complete_frame(c);
        if (rc != SRB_OK) return rc;
    }
    return launch_frame(c);
}

#ifdef SRB_TIMELINE
// Diagnostic builds only: the last min(frames launched, max_frames) frames' stamps, 4 x uint64 ns each
// (geometry start, geometry end, tile kernel start, tile kernel end); returns the number of frames written.
int srb_debug_timeline(SrbContext* c, unsigned long long* out, int max_frames) {
    if (!c || !out) return SRB_ERR_ARG;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    const uint32_t n = std::min<uint32_t>(std::min<uint32_t>(c->timelineFrame, kTimelineFrames), (uint32_t)std::max(max_frames, 0));
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t f = c->timelineFrame - n + i;
        cudaMemcpy(out + 4 * (size_t)i, c->dTimeline + 4 * (size_t)(f % kTimelineFrames), 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    }
    return (int)n;
}
#endif

int srb_sync(SrbContext* c) {
    if (!c) return SRB_ERR_ARG;
    CK(c, cudaSetDevice(c->device));
    return complete_frame(c);
}

int srb_read_color_rows(SrbContext* c, float* out, int y0, int y1) {
    if (!c || !out) return SRB_ERR_ARG;
    if (y0 < 0 || y1 < y0 || y1 > c->H) return fail(c, SRB_ERR_ARG, "row range out of bounds");
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    size_t off = (size_t)y0 * c->W, n = (size_t)(y1 - y0) * c->W;
    CK(c, cudaMemcpyAsync(out, c->color + off, n * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    return SRB_OK;
}

int srb_read_color(SrbContext* c, float* out) { return c ? srb_read_color_rows(c, out, 0, c->H) : SRB_ERR_ARG; }

int srb_read_depth(SrbContext* c, float* out) {
    if (!c || !out) return SRB_ERR_ARG;
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    CK(c, cudaMemcpyAsync(out, c->depth, (size_t)c->W * c->H * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    return SRB_OK;
}

static int pixel_index(SrbContext* c, int x, int y, size_t* idx) {
    if (x < 0 || x >= c->W || y < 0 || y >= c->H) return 0;     // accessors ignore out-of-range (MainWindow.cs:384,392)
    *idx = (size_t)y * c->W + x;
    return 1;
}
int srb_get_pixel(SrbContext* c, int x, int y, float out[4]) {
    if (!c || !out) return SRB_ERR_ARG;
    size_t i;
    if (!pixel_index(c, x, y, &i)) { out[0] = out[1] = out[2] = out[3] = 0.0f; return SRB_OK; }   // Vector4.Zero (:397)
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    CK(c, cudaMemcpy(out, c->color + i, 16, cudaMemcpyDeviceToHost));
    return SRB_OK;
}
int srb_set_pixel(SrbContext* c, int x, int y, const float rgba[4]) {
    if (!c || !rgba) return SRB_ERR_ARG;
    size_t i;
    if (!pixel_index(c, x, y, &i)) return SRB_OK;
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    CK(c, cudaMemcpy(c->color + i, rgba, 16, cudaMemcpyHostToDevice));
    return SRB_OK;
}
int srb_get_depth(SrbContext* c, int x, int y, float* out) {
    if (!c || !out) return SRB_ERR_ARG;
    size_t i;
    if (!pixel_index(c, x, y, &i)) { *out = kClearDepth; return SRB_OK; }     // float.MinValue (:426)
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    CK(c, cudaMemcpy(out, c->depth + i, 4, cudaMemcpyDeviceToHost));
    return SRB_OK;
}
int srb_set_depth(SrbContext* c, int x, int y, float depth) {
    if (!c) return SRB_ERR_ARG;
    size_t i;
    if (!pixel_index(c, x, y, &i)) return SRB_OK;
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    CK(c, cudaMemcpy(c->depth + i, &depth, 4, cudaMemcpyHostToDevice));
    return SRB_OK;
}

int srb_read_color_rgba8(SrbContext* c, uint8_t* out) {
    if (!c || !out) return SRB_ERR_ARG;
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    size_t n = (size_t)c->W * c->H;
    CK(c, c->dBinScratch.reserve(n));            // reuse scratch: 4 B per pixel
    pack_rgba8_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->color, (uchar4*)c->dBinScratch.p, n);
    CK(c, cudaGetLastError());
    CK(c, cudaMemcpyAsync(out, c->dBinScratch.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    return SRB_OK;
}

/* Present path (MainWindow.cs:226-246).  Flatten on the device, copy on a second stream: the call returns
 * once the copy is queued, and the next frame may be submitted and rendered while it runs. */
// gathered: the source is this context's RGBA8 gather frame (already 8-bit, assembled by all band ranks): staged
// with a device copy instead of a pack kernel, so that the next frame may overwrite the gather frame at once.
static int present_rows_async(SrbContext* c, void* out, int y0, int y1, bool rgba8, bool gathered = false) {
    if (!c || !out) return SRB_ERR_ARG;
    if (gathered && !c->color8) return fail(c, SRB_ERR_STATE, "no RGBA8 gather frame on this context (srb_gather_rgba8_export first)");
    if (y0 < 0 || y1 < y0 || y1 > c->H) return fail(c, SRB_ERR_ARG, "row range out of bounds");
    CK(c, cudaSetDevice(c->device));
    int rc = complete_frame(c, false);                     // capacity overflows settled; the flatten below is queued
    if (rc != SRB_OK) return rc;                           // behind the frame's tile kernel, nobody waits for it here
    if (y1 == y0) return SRB_OK;
    const size_t n = (size_t)c->W * (y1 - y0), first = (size_t)c->W * y0;
    if (!c->copyStream) CK(c, cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking));
    const int s = c->presentSlot;
    c->presentSlot ^= 1;
    if (!c->presentReady[s]) {
        CK(c, cudaEventCreateWithFlags(&c->presentReady[s], cudaEventDisableTiming));
        CK(c, cudaEventCreateWithFlags(&c->presentCopied[s], cudaEventDisableTiming));
    }
    if (c->presentBusy[s]) {                               // the copy that used this slot two frames ago
        CK(c, cudaEventSynchronize(c->presentCopied[s]));
        c->presentBusy[s] = false;
    }
    CK(c, c->dPresent[s].reserve((rgba8 ? n : n * 3) + 16));
    if (gathered)
        CK(c, cudaMemcpyAsync(c->dPresent[s].p, c->color8 + first, n * 4, cudaMemcpyDeviceToDevice, c->stream));
    else if (rgba8)
        present_rgba8_kernel<<<(unsigned)((n / 4 + 1 + 255) / 256), 256, 0, c->stream>>>(c->color + first, (uint32_t*)c->dPresent[s].p, n);
    else
        present_rgb_kernel<<<(unsigned)((n / 4 + 1 + 255) / 256), 256, 0, c->stream>>>(c->color + first, c->dPresent[s].p, n);
    CK(c, cudaGetLastError());
    CK(c, cudaEventRecord(c->presentReady[s], c->stream));
    CK(c, cudaStreamWaitEvent(c->copyStream, c->presentReady[s], 0));
    // (Measured, no gain: the copy as two halves on two streams -- one stream already reaches the link rate, 54.7 GB/s.)
    if (rgba8)
        CK(c, cudaMemcpyAsync((uint8_t*)out + first * 4, c->dPresent[s].p, n * 4, cudaMemcpyDeviceToHost, c->copyStream));
    else
        CK(c, cudaMemcpyAsync((float*)out + first * 3, c->dPresent[s].p, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, c->copyStream));
    CK(c, cudaEventRecord(c->presentCopied[s], c->copyStream));
    c->presentBusy[s] = true;
    return SRB_OK;
}

int srb_present_rgb_rows_async(SrbContext* c, float* out, int y0, int y1) { return present_rows_async(c, out, y0, y1, false); }
int srb_present_rgba8_rows_async(SrbContext* c, uint8_t* out, int y0, int y1) { return present_rows_async(c, out, y0, y1, true); }
int srb_present_rgba8_async(SrbContext* c, uint8_t* out) {
    return c ? present_rows_async(c, out, 0, c->H, true) : SRB_ERR_ARG;
}
int srb_present_gathered_rgba8_async(SrbContext* c, uint8_t* out) {
    return c ? present_rows_async(c, out, 0, c->H, true, true) : SRB_ERR_ARG;
}
int srb_present_rgba8(SrbContext* c, uint8_t* out) {
    int rc = srb_present_rgba8_async(c, out);
    return rc != SRB_OK ? rc : srb_present_wait(c);
}

int srb_present_rgb_async(SrbContext* c, float* out) {
    return c ? srb_present_rgb_rows_async(c, out, 0, c->H) : SRB_ERR_ARG;
}

int srb_present_wait(SrbContext* c) {
    if (!c) return SRB_ERR_ARG;
    CK(c, cudaSetDevice(c->device));
    for (int s = 0; s < 2; ++s)
        if (c->presentBusy[s]) {
            CK(c, cudaEventSynchronize(c->presentCopied[s]));
            c->presentBusy[s] = false;
        }
    return SRB_OK;
}

int srb_present_fence(SrbContext* c) {
    if (!c) return SRB_ERR_ARG;
    CK(c, cudaSetDevice(c->device));
    for (int s = 0; s < 2; ++s)
        if (c->presentBusy[s]) CK(c, cudaStreamWaitEvent(c->stream, c->presentCopied[s], 0));
    return SRB_OK;
}

/* ---- peer-memory frame barrier ---- */
int srb_barrier_export(SrbContext* c, uint8_t handle_out[SRB_IPC_HANDLE_BYTES]) {
    if (!c || !handle_out) return SRB_ERR_ARG;
    CK(c, cudaSetDevice(c->device));
    int rc = complete_frame(c);
    if (rc != SRB_OK) return rc;
    if (!c->barRow) {
        CK(c, cudaMalloc((void**)&c->barRow, (size_t)kMaxRanks * kFlagStride * sizeof(uint32_t)));
        CK(c, cudaMallocHost((void**)&c->barTimedOut, sizeof(uint32_t)));
    }
    CK(c, cudaMemset(c->barRow, 0, (size_t)kMaxRanks * kFlagStride * sizeof(uint32_t)));
    *c->barTimedOut = 0u;
    c->barCount = 0;
    cudaIpcMemHandle_t h;
    CK(c, cudaIpcGetMemHandle(&h, c->barRow));
    std::memcpy(handle_out, &h, sizeof(h));
    return SRB_OK;
}

int srb_barrier_connect(SrbContext* c, int rank, int world, const uint8_t* handles) {
    if (!c) return SRB_ERR_ARG;
    CK(c, cudaSetDevice(c->device));
    int rc = complete_frame(c);
    if (rc != SRB_OK) return rc;
    barrier_close(c);
    if (!handles) return SRB_OK;                                  // disconnect
    if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return fail(c, SRB_ERR_ARG, "barrier: need 0 <= rank < world <= 16");
    if (!c->barRow) return fail(c, SRB_ERR_STATE, "srb_barrier_export first");
    for (int r = 0; r < world; ++r) {
        if (r == rank) { c->barRows.row[r] = c->barRow; continue; }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + (size_t)r * SRB_IPC_HANDLE_BYTES, sizeof(h));
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            c->barWorld = r; c->barRank = rank;
            barrier_close(c);
            return fail(c, SRB_ERR_CUDA, std::string("cudaIpcOpenMemHandle (barrier): ") + cudaGetErrorString(e));
        }
        c->barRows.row[r] = (uint32_t*)p;
    }
    c->barRank = rank; c->barWorld = world;
    return SRB_OK;
}

int srb_frame_barrier(SrbContext* c, int slack) {
    if (!c) return SRB_ERR_ARG;
    if (c->barWorld <= 0) return fail(c, SRB_ERR_STATE, "frame barrier is not connected");
    if (slack < 0 || slack > 2) return fail(c, SRB_ERR_ARG, "barrier slack must be 0..2 frames");
    if (*c->barTimedOut) return fail(c, SRB_ERR_CUDA, "frame barrier timed out: a rank is missing or far behind");
    CK(c, cudaSetDevice(c->device));
    const uint32_t k = ++c->barCount;
    const uint32_t waitFor = k > (uint32_t)slack ? k - (uint32_t)slack : 0u;
    if (!c->barStream) {
        CK(c, cudaStreamCreateWithFlags(&c->barStream, cudaStreamNonBlocking));
        CK(c, cudaEventCreateWithFlags(&c->barFrame, cudaEventDisableTiming));
        for (auto& e : c->barDone) CK(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    if (slack == 0) {
        // The caller reads the assembled frame next: the barrier sits on the raster stream itself (behind any
        // side-stream barrier still in flight, so that this rank's counter never steps back).
        if (c->barPending >= 0) { CK(c, cudaStreamWaitEvent(c->stream, c->barDone[c->barPending], 0)); c->barPending = -1; }
        frame_barrier_kernel<<<1, 32, 0, c->stream>>>(c->barRows, c->barRank, c->barWorld, k, waitFor, c->barTimedOut);
    } else {
        // Ranks only have to stay close: the kernel runs on a stream of its own behind this frame's tile kernel, and
        // the raster stream waits for the PREVIOUS call's barrier -- in effect slack + 1 frames.  (On the raster
        // stream, between tile kernels k and k+1, the launch gap and the NVLink round trip of the flag stores cost
        // 5 % of an N=2 city4k frame: 7,400 -> 7,050 frames/s.)
        CK(c, cudaEventRecord(c->barFrame, c->stream));
        CK(c, cudaStreamWaitEvent(c->barStream, c->barFrame, 0));
        frame_barrier_kernel<<<1, 32, 0, c->barStream>>>(c->barRows, c->barRank, c->barWorld, k, waitFor, c->barTimedOut);
        const int cur = c->barPending == 0 ? 1 : 0;
        CK(c, cudaEventRecord(c->barDone[cur], c->barStream));
        if (c->barPending >= 0) CK(c, cudaStreamWaitEvent(c->stream, c->barDone[c->barPending], 0));
        c->barPending = cur;
    }
    CK(c, cudaGetLastError());
    return SRB_OK;
}

int srb_present_rgb(SrbContext* c, float* out) {
    int rc = srb_present_rgb_async(c, out);
    return rc != SRB_OK ? rc : srb_present_wait(c);
}

int srb_frustum_test(SrbContext* c, const float* spheres, const float* models, const float view[16],
                     const float proj[16], int n, int32_t* out_visible) {
    if (!c || !spheres || !models || !view || !proj || !out_visible || n < 0) return SRB_ERR_ARG;
    if (n == 0) return SRB_OK;
    CK(c, cudaSetDevice(c->device));
    float* d = nullptr;
    size_t fl = (size_t)n * 20 + 32;
    CK(c, cudaMalloc((void**)&d, fl * sizeof(float) + (size_t)n * sizeof(int32_t)));
    float* dS = d; float* dM = d + (size_t)n * 4; float* dVP = dM + (size_t)n * 16;
    int32_t* dOut = (int32_t*)(dVP + 32);
    cudaError_t e = cudaMemcpyAsync(dS, spheres, (size_t)n * 16, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dM, models, (size_t)n * 64, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dVP, view, 64, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dVP + 16, proj, 64, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        frustum_batch_kernel<<<(n + 127) / 128, 128, 0, c->stream>>>(dS, dM, dVP, n, c->uniforms.numerics_fma, dOut);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_visible, dOut, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if (e != cudaSuccess) return fail(c, SRB_ERR_CUDA, std::string("srb_frustum_test: ") + cudaGetErrorString(e));
    return SRB_OK;
}

// Matrix4x4.Invert of the .NET BCL (scalar algorithm: cofactors of the first column, fails when |det| < float.Epsilon),
// restated from its published form; row-major M11..M44.  Only the hit normal depends on it.
static bool invert4(const float* M, float* R) {
    const float a = M[0], b = M[1], c = M[2], d = M[3], e = M[4], f = M[5], g = M[6], h = M[7];
    const float i = M[8], j = M[9], k = M[10], l = M[11], m = M[12], n = M[13], o = M[14], p = M[15];
    const float kp_lo = k * p - l * o, jp_ln = j * p - l * n, jo_kn = j * o - k * n;
    const float ip_lm = i * p - l * m, io_km = i * o - k * m, in_jm = i * n - j * m;
    const float a11 = +(f * kp_lo - g * jp_ln + h * jo_kn), a12 = -(e * kp_lo - g * ip_lm + h * io_km);
    const float a13 = +(e * jp_ln - f * ip_lm + h * in_jm), a14 = -(e * jo_kn - f * io_km + g * in_jm);
    const float det = a * a11 + b * a12 + c * a13 + d * a14;
    if (std::fabs(det) < 1.401298464e-45f) return false;
    const float invDet = 1.0f / det;
    R[0] = a11 * invDet; R[4] = a12 * invDet; R[8] = a13 * invDet; R[12] = a14 * invDet;
    R[1] = -(b * kp_lo - c * jp_ln + d * jo_kn) * invDet; R[5] = +(a * kp_lo - c * ip_lm + d * io_km) * invDet;
    R[9] = -(a * jp_ln - b * ip_lm + d * in_jm) * invDet; R[13] = +(a * jo_kn - b * io_km + c * in_jm) * invDet;
    const float gp_ho = g * p - h * o, fp_hn = f * p - h * n, fo_gn = f * o - g * n;
    const float ep_hm = e * p - h * m, eo_gm = e * o - g * m, en_fm = e * n - f * m;
    R[2] = +(b * gp_ho - c * fp_hn + d * fo_gn) * invDet; R[6] = -(a * gp_ho - c * ep_hm + d * eo_gm) * invDet;
    R[10] = +(a * fp_hn - b * ep_hm + d * en_fm) * invDet; R[14] = -(a * fo_gn - b * eo_gm + c * en_fm) * invDet;
    const float gl_hk = g * l - h * k, fl_hj = f * l - h * j, fk_gj = f * k - g * j;
    const float el_hi = e * l - h * i, ek_gi = e * k - g * i, ej_fi = e * j - f * i;
    R[3] = -(b * gl_hk - c * fl_hj + d * fk_gj) * invDet; R[7] = +(a * gl_hk - c * el_hi + d * ek_gi) * invDet;
    R[11] = -(a * fl_hj - b * el_hi + d * ej_fi) * invDet; R[15] = +(a * fk_gj - b * ek_gi + c * ej_fi) * invDet;
    return true;
}

/* Physics.Raycast (Physics.cs:19-179), batched: n rays against one resident mesh. */
int srb_raycast(SrbContext* c, SrbMesh mesh, const float model[16], const float* rays, int n, int face_mask, SrbRayHit* out) {
    static_assert(sizeof(SrbRayHit) == sizeof(RayHit), "SrbRayHit layout");
    if (!c || !model || !rays || !out || n < 0) return SRB_ERR_ARG;
    if (mesh <= 0 || (size_t)mesh > c->meshes.size() || !c->meshes[mesh - 1].alive) return fail(c, SRB_ERR_HANDLE, "unknown mesh handle");
    if (face_mask < 0 || face_mask > 3) return fail(c, SRB_ERR_ARG, "face_mask must be a combination of 1 (IgnoreBackfaces) and 2 (IgnoreFrontfaces)");
    if (n == 0) return SRB_OK;
    CK(c, cudaSetDevice(c->device));
    const MeshRes& m = c->meshes[mesh - 1];
    Mat4 M, NM;
    std::memcpy(M.m, model, 64);
    float inv[16];
    if (!invert4(model, inv)) {                                  // :30-36: no hit, distance float.MaxValue, zero vectors
        for (int r = 0; r < n; ++r) {
            std::memset(&out[r], 0, sizeof(SrbRayHit));
            out[r].distance = 3.402823466e+38f; out[r].triangle = -1;
        }
        return SRB_OK;
    }
    for (int r = 0; r < 4; ++r) for (int k = 0; k < 4; ++k) NM.m[4 * r + k] = inv[4 * k + r];     // Matrix4x4.Transpose (:38)
    // scratch: rays (6 floats), normalised directions (float4), keys (u64), hits
    const size_t bytes = (size_t)n * (24 + 16 + 8 + sizeof(RayHit)) + 64;
    CK(c, c->dRayScratch.reserve((bytes + 15) / 16));
    char* base = (char*)c->dRayScratch.p;
    float4* dDirs = (float4*)base;
    RayHit* dHits = (RayHit*)(base + (size_t)n * 16);
    unsigned long long* dKeys = (unsigned long long*)(base + (size_t)n * (16 + sizeof(RayHit)));
    float* dRays = (float*)(base + (size_t)n * (16 + sizeof(RayHit) + 8));
    cudaStream_t s = c->stream;
    CK(c, cudaMemcpyAsync(dRays, rays, (size_t)n * 24, cudaMemcpyHostToDevice, s));
    ray_prepare_kernel<<<(n + 255) / 256, 256, 0, s>>>(dRays, n, dDirs, dKeys);
    const int nTri = m.nI / 3;
    if (nTri > 0) {
        dim3 grid((nTri + 255) / 256, (n + kRayChunk - 1) / kRayChunk);
        if (grid.y > 65535) return fail(c, SRB_ERR_ARG, "too many rays in one call (max 65535 * 64)");
        if (c->uniforms.numerics_fma) raycast_kernel<true><<<grid, 256, 0, s>>>(m.vtx, m.idx, nTri, M, dRays, dDirs, n, face_mask, dKeys);
        else raycast_kernel<false><<<grid, 256, 0, s>>>(m.vtx, m.idx, nTri, M, dRays, dDirs, n, face_mask, dKeys);
    }
    if (c->uniforms.numerics_fma) ray_resolve_kernel<true><<<(n + 127) / 128, 128, 0, s>>>(m.vtx, m.idx, M, NM, dRays, dDirs, n, face_mask, dKeys, dHits);
    else ray_resolve_kernel<false><<<(n + 127) / 128, 128, 0, s>>>(m.vtx, m.idx, M, NM, dRays, dDirs, n, face_mask, dKeys, dHits);
    CK(c, cudaGetLastError());
    CK(c, cudaMemcpyAsync(out, dHits, (size_t)n * sizeof(RayHit), cudaMemcpyDeviceToHost, s));
    CK(c, cudaStreamSynchronize(s));
    return SRB_OK;
}

/* ---- srb_raycast without any allocation ----
 * No counterpart in the reference; added for callers that cannot allocate (CUDA
 * graph capture, tight serving loops).  Same kernels as srb_raycast, the per-call scratch (ray copies, normalised
 * directions, 64-bit winner keys, hit records: 96 bytes per ray) lives in ONE file-scope device array of fixed size
 * instead of the context's growable buffer.  Every GPU that loads the library holds the array until the process ends.
 * All calls share it, whatever the context: it keeps the last call's contents (ray_prepare_kernel rewrites every key
 * and direction a call uses before they are read), and calls that overlap on the GPU would race -- this entry point
 * serialises on the context's stream and waits for its result, so only concurrent callers on DIFFERENT contexts of
 * one device have to take turns themselves. */
namespace {
constexpr size_t kStaticRayBytes = (size_t)SRB_RAYCAST_STATIC_MAX_RAYS * (24 + 16 + 8 + sizeof(RayHit));
static __device__ __align__(256) unsigned char g_rayScratch[kStaticRayBytes];
std::mutex g_rayScratchMu;
unsigned char* g_rayScratchAddr[64] = {};                       // per device: the array's DEVICE address (the name above is
                                                                // only a host-side stand-in and must never reach a kernel)
int static_scratch_address(SrbContext* c, unsigned char** out) {
    if (c->device < 0 || c->device >= 64) return fail(c, SRB_ERR_ARG, "device index beyond the static scratch table");
    std::lock_guard<std::mutex> lk(g_rayScratchMu);
    if (!g_rayScratchAddr[c->device]) {                         // once per device, with that device current
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_rayScratch);
        if (e != cudaSuccess || !p) return fail(c, SRB_ERR_CUDA, std::string("cudaGetSymbolAddress(g_rayScratch): ") + cudaGetErrorString(e));
        g_rayScratchAddr[c->device] = (unsigned char*)p;
    }
    *out = g_rayScratchAddr[c->device];
    return SRB_OK;
}
}  // namespace

int srb_raycast_static(SrbContext* c, SrbMesh mesh, const float model[16], const float* rays, int n, int face_mask, SrbRayHit* out) {
    if (!c || !model || !rays || !out || n < 0) return SRB_ERR_ARG;
    if (n > SRB_RAYCAST_STATIC_MAX_RAYS)
        return fail(c, SRB_ERR_ARG, "srb_raycast_static: at most " + std::to_string(SRB_RAYCAST_STATIC_MAX_RAYS) +
                                    " rays per call (fixed device scratch); use srb_raycast or split the batch");
    if (mesh <= 0 || (size_t)mesh > c->meshes.size() || !c->meshes[mesh - 1].alive) return fail(c, SRB_ERR_HANDLE, "unknown mesh handle");
    if (face_mask < 0 || face_mask > 3) return fail(c, SRB_ERR_ARG, "face_mask must be a combination of 1 (IgnoreBackfaces) and 2 (IgnoreFrontfaces)");
    if (n == 0) return SRB_OK;
    CK(c, cudaSetDevice(c->device));
    const MeshRes& m = c->meshes[mesh - 1];
    Mat4 M, NM;
    std::memcpy(M.m, model, 64);
    float inv[16];
    if (!invert4(model, inv)) {                                  // Physics.cs:30-36: no hit, distance float.MaxValue, zero vectors
        for (int r = 0; r < n; ++r) {
            std::memset(&out[r], 0, sizeof(SrbRayHit));
            out[r].distance = 3.402823466e+38f; out[r].triangle = -1;
        }
        return SRB_OK;
    }
    for (int r = 0; r < 4; ++r) for (int k = 0; k < 4; ++k) NM.m[4 * r + k] = inv[4 * k + r];
    unsigned char* base = nullptr;
    int rc = static_scratch_address(c, &base);
    if (rc != SRB_OK) return rc;
    // fixed layout for the maximum batch: directions | hits | keys | rays (each region 16-byte aligned)
    const size_t N = SRB_RAYCAST_STATIC_MAX_RAYS;
    float4* dDirs = (float4*)base;
    RayHit* dHits = (RayHit*)(base + N * 16);
    unsigned long long* dKeys = (unsigned long long*)(base + N * (16 + sizeof(RayHit)));
    float* dRays = (float*)(base + N * (16 + sizeof(RayHit) + 8));
    cudaStream_t s = c->stream;
    CK(c, cudaMemcpyAsync(dRays, rays, (size_t)n * 24, cudaMemcpyHostToDevice, s));
    ray_prepare_kernel<<<(n + 255) / 256, 256, 0, s>>>(dRays, n, dDirs, dKeys);      // (re)initialises every key and direction used below
    const int nTri = m.nI / 3;
    if (nTri > 0) {
        dim3 grid((nTri + 255) / 256, (n + kRayChunk - 1) / kRayChunk);
        if (c->uniforms.numerics_fma) raycast_kernel<true><<<grid, 256, 0, s>>>(m.vtx, m.idx, nTri, M, dRays, dDirs, n, face_mask, dKeys);
        else raycast_kernel<false><<<grid, 256, 0, s>>>(m.vtx, m.idx, nTri, M, dRays, dDirs, n, face_mask, dKeys);
    }
    if (c->uniforms.numerics_fma) ray_resolve_kernel<true><<<(n + 127) / 128, 128, 0, s>>>(m.vtx, m.idx, M, NM, dRays, dDirs, n, face_mask, dKeys, dHits);
    else ray_resolve_kernel<false><<<(n + 127) / 128, 128, 0, s>>>(m.vtx, m.idx, M, NM, dRays, dDirs, n, face_mask, dKeys, dHits);
    CK(c, cudaGetLastError());
    CK(c, cudaMemcpyAsync(out, dHits, (size_t)n * sizeof(RayHit), cudaMemcpyDeviceToHost, s));
    CK(c, cudaStreamSynchronize(s));
    return SRB_OK;
}

void* srb_device_color_ptr(SrbContext* c) { return c ? (void*)c->color : nullptr; }
void* srb_device_depth_ptr(SrbContext* c) { return c ? (void*)c->depth : nullptr; }
void* srb_stream(SrbContext* c) { return c ? (void*)c->stream : nullptr; }

int srb_set_profiling(SrbContext* c, int enabled) {
    if (!c) return SRB_ERR_ARG;
    CK(c, cudaSetDevice(c->device));
    int rc = complete_frame(c);
    if (rc != SRB_OK) return rc;
    if (enabled && !c->stageEv[0])
        for (int i = 0; i <= SRB_N_STAGES; ++i) CK(c, cudaEventCreate(&c->stageEv[i]));
    c->profiling = enabled != 0;
    return SRB_OK;
}

int srb_get_stage_ms(SrbContext* c, float* out, int n) {
    if (!c || !out || n < 0) return SRB_ERR_ARG;
    if (!c->profiling) return fail(c, SRB_ERR_STATE, "profiling is off");
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    for (int i = 0; i < n && i < SRB_N_STAGES; ++i) out[i] = c->stageMs[i];
    return SRB_OK;
}

#ifdef SRB_TILE_TIMING
/* Diagnostic builds only (not declared in srb200.h): per-CTA timing of the last visibility tile kernel. */
int srb_debug_tile_times(SrbContext* c, unsigned long long* out, int nTiles) {
    if (!c || !out) return SRB_ERR_ARG;
    int rc = srb_sync(c);
    if (rc != SRB_OK) return rc;
    CK(c, cudaMemcpy(out, c->dBinScratch.p, (size_t)nTiles * 32, cudaMemcpyDeviceToHost));
    return SRB_OK;
}
#endif

int srb_get_stats(SrbContext* c, SrbStats* out) {
    if (!c || !out) return SRB_ERR_ARG;
    *out = c->stats;
    return SRB_OK;
}

}  // extern "C"
