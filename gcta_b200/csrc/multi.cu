// multi.cu - gcta_b200_make_grm_multi: the whole --make-grm path on several GPUs of ONE process.
//
// The reference calls make_grm once per process from the main thread (option.cpp:620-621); this is that call for a
// multi-GPU box.  The lower triangle is cut into row bands (gcta_b200_partition), one plan and one host thread per
// device, no data-path collective.  What every device needs is the packed genotype file: instead of one full upload
// per device (G x the file over PCIe), the file rows are cut into the exchange's segments (gcta_b200_exchange_slice);
// of each segment device g uploads 1/G over its own PCIe link and pushes that slice into the other devices' buffers
// with peer copies (copy engines over NVLink; no SM, so nothing competes with the persistent SYRK grid).  A band starts
// SNP chunk c as soon as all G slices of the last segment the chunk reads have been pushed: per device one "poster"
// thread feeds the copies, the band's own thread runs the plan and is held back per chunk by a call-back
// (gcta_b200_plan_set_chunk_hook) that waits - on the host for the event to exist, on the device for it to fire.
// Host buffers may be pageable: the uploads then block their poster thread only.
#include <algorithm>
#include <atomic>
#include <memory>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

struct Shared {
    int G = 1;
    int64_t nseg = 0, m_file = 0, pitch = 0, stride = 0;
    std::vector<gcta_b200_plan*> plans;
    std::vector<uint8_t*> beds;                       // plan g's genotype buffer
    std::vector<int> devs;
    std::vector<std::vector<cudaEvent_t>> pushed;     // [g][s]: device g's slice of segment s has landed on every device
    std::unique_ptr<std::atomic<int64_t>[]> posted;   // [g]: number of segments whose `pushed` event has been recorded
    std::atomic<int> failed{0};
};

struct HookCtx {
    Shared* sh;
    std::vector<int64_t> last_seg;   // per SNP chunk of this band's plan: the last segment it reads
};

int chunk_hook(void* vctx, int64_t c, void* stream)
{
    HookCtx* h = (HookCtx*)vctx;
    Shared* sh = h->sh;
    const int64_t s = h->last_seg[(size_t)c];
    for (int q = 0; q < sh->G; q++) {
        // a wait captures the event's latest record, so the record must exist before the wait is queued
        while (sh->posted[q].load(std::memory_order_acquire) <= s) {
            if (sh->failed.load()) return 1;
            std::this_thread::yield();
        }
        // the pushes of one device run in order on one stream: its segment s implies all earlier ones
        if (cudaStreamWaitEvent((cudaStream_t)stream, sh->pushed[q][(size_t)s], 0) != cudaSuccess) return 1;
    }
    return 0;
}

// upload this device's slice of every segment and push it to the peers
void post_all(Shared* sh, int g, const uint8_t* bed_host)
{
    cudaStream_t s_up = nullptr, s_push = nullptr;
    std::vector<cudaEvent_t> up((size_t)sh->nseg, nullptr);
    bool ok = cudaSetDevice(sh->devs[g]) == cudaSuccess && cudaStreamCreateWithFlags(&s_up, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&s_push, cudaStreamNonBlocking) == cudaSuccess;
    for (int64_t s = 0; ok && s < sh->nseg; s++) {
        int64_t r0 = 0, r1 = 0;
        gcta_b200_exchange_slice(sh->m_file, sh->G, g, s, &r0, &r1);
        if (r1 > r0) {
            ok = cudaMemcpy2DAsync(sh->beds[g] + r0 * sh->pitch, (size_t)sh->pitch, bed_host + r0 * sh->stride, (size_t)sh->stride,
                                   (size_t)sh->stride, (size_t)(r1 - r0), cudaMemcpyHostToDevice, s_up) == cudaSuccess;
            ok = ok && cudaEventCreateWithFlags(&up[(size_t)s], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventRecord(up[(size_t)s], s_up) == cudaSuccess && cudaStreamWaitEvent(s_push, up[(size_t)s], 0) == cudaSuccess;
            for (int q = 0; ok && q < sh->G; q++)
                if (q != g)
                    ok = cudaMemcpyPeerAsync(sh->beds[q] + r0 * sh->pitch, sh->devs[q], sh->beds[g] + r0 * sh->pitch, sh->devs[g],
                                             (size_t)(r1 - r0) * sh->pitch, s_push) == cudaSuccess;
        }
        ok = ok && cudaEventRecord(sh->pushed[g][(size_t)s], s_push) == cudaSuccess;
        if (ok) sh->posted[g].store(s + 1, std::memory_order_release);
    }
    if (!ok) sh->failed.store(1);
    if (s_up) cudaStreamSynchronize(s_up);
    if (s_push) cudaStreamSynchronize(s_push);
    for (cudaEvent_t e : up)
        if (e) cudaEventDestroy(e);
    if (s_up) cudaStreamDestroy(s_up);
    if (s_push) cudaStreamDestroy(s_push);
    cudaGetLastError();
}

}  // namespace

extern "C" int gcta_b200_make_grm_multi(const gcta_b200_args* a, int32_t n_gpus, gcta_b200_out* out, gcta_b200_out* per_gpu,
                                        char* err, size_t errcap)
{
    using namespace gb;
    GB_ARG(a && out && a->bed && out->A_lower, "make_grm_multi: null argument");
    GB_ARG(n_gpus >= 1, "make_grm_multi: n_gpus must be at least 1");
    const int ndev = gcta_b200_device_count();
    if (ndev <= 0) return set_err(err, errcap, GCTA_B200_ERR_NO_DEVICE, "Error: no CUDA device is available (gcta_b200 has no CPU fallback).");
    GB_ARG(a->device >= 0 && a->device < ndev, "make_grm_multi: bad device ordinal");
    const int G = std::min<int>(n_gpus, ndev - a->device);
    if (G == 1) {
        gcta_b200_args a1 = *a;
        a1.part_rank = 0;
        a1.part_world = 1;
        int rc = gcta_b200_make_grm(&a1, out, err, errcap);
        if (!rc && per_gpu) per_gpu[0] = *out;
        return rc;
    }

    Shared sh;
    sh.G = G;
    sh.plans.assign(G, nullptr);
    sh.beds.assign(G, nullptr);
    sh.devs.resize(G);
    sh.posted.reset(new std::atomic<int64_t>[G]);
    for (int g = 0; g < G; g++) { sh.devs[g] = a->device + g; sh.posted[g].store(0); }
    std::vector<std::string> errs(G);
    std::vector<int> rcs(G, 0);
    std::vector<gcta_b200_out> outs(G);
    auto on_each = [&](auto&& fn) {   // one host thread per device
        std::vector<std::thread> th;
        for (int g = 0; g < G; g++) th.emplace_back(fn, g);
        for (auto& t : th) t.join();
    };
    auto cleanup = [&]() {
        for (int g = 0; g < G; g++) {
            if (sh.plans[g]) gcta_b200_plan_destroy(sh.plans[g]);
            if (g < (int)sh.pushed.size()) {
                cudaSetDevice(sh.devs[g]);
                for (cudaEvent_t e : sh.pushed[g])
                    if (e) cudaEventDestroy(e);
            }
        }
        cudaGetLastError();
    };
    auto first_error = [&]() -> int {
        for (int g = 0; g < G; g++)
            if (rcs[g]) return set_err(err, errcap, rcs[g], errs[g]);
        return 0;
    };

    // one plan per device (the allocations of the devices run side by side)
    on_each([&](int g) {
        gcta_b200_args ag = *a;
        ag.device = sh.devs[g];
        ag.part_rank = g;
        ag.part_world = G;
        char e[512] = {0};
        rcs[g] = gcta_b200_plan_create(&ag, &sh.plans[g], e, sizeof e);
        errs[g] = e;
    });
    if (int rc = first_error()) { cleanup(); return rc; }
    gcta_b200_plan_geometry(sh.plans[0], nullptr, &sh.m_file, &sh.stride, &sh.pitch);
    sh.nseg = gcta_b200_exchange_slice(sh.m_file, G, 0, 0, nullptr, nullptr);
    bool resident = true;
    for (int g = 0; g < G; g++) {
        sh.beds[g] = (uint8_t*)gcta_b200_plan_own_bed(sh.plans[g]);
        resident = resident && sh.beds[g] != nullptr;
    }

    std::vector<HookCtx> hooks(G);
    if (resident) {
        // direct peer copies where the devices allow it (otherwise cudaMemcpyPeerAsync stages through the host)
        for (int g = 0; g < G; g++) {
            cudaSetDevice(sh.devs[g]);
            for (int q = 0; q < G; q++)
                if (q != g) {
                    int can = 0;
                    if (cudaDeviceCanAccessPeer(&can, sh.devs[g], sh.devs[q]) == cudaSuccess && can) cudaDeviceEnablePeerAccess(sh.devs[q], 0);
                }
            cudaGetLastError();   // "already enabled" is fine
        }
        sh.pushed.assign(G, std::vector<cudaEvent_t>((size_t)sh.nseg, nullptr));
        for (int g = 0; g < G; g++) {
            cudaSetDevice(sh.devs[g]);
            for (int64_t s = 0; s < sh.nseg; s++)
                if (cudaEventCreateWithFlags(&sh.pushed[g][(size_t)s], cudaEventDisableTiming) != cudaSuccess) {
                    cleanup();
                    return set_err(err, errcap, GCTA_B200_ERR_CUDA, "Error: make_grm_multi: cannot create the exchange events.");
                }
            int64_t s0 = 0, s1 = 0;
            gcta_b200_exchange_slice(sh.m_file, 1, 0, 0, &s0, &s1);   // with one rank a slice is the whole segment
            const int64_t rows_per_seg = std::max<int64_t>(1, s1 - s0);
            hooks[g].sh = &sh;
            const int64_t nch = gcta_b200_plan_num_chunks(sh.plans[g]);
            hooks[g].last_seg.resize((size_t)nch);
            for (int64_t c = 0; c < nch; c++) {
                int64_t r0 = 0, r1 = 0;
                gcta_b200_plan_chunk_rows(sh.plans[g], c, &r0, &r1);
                hooks[g].last_seg[(size_t)c] = std::min<int64_t>(sh.nseg - 1, (std::max<int64_t>(r1, 1) - 1) / rows_per_seg);
            }
            gcta_b200_plan_set_chunk_hook(sh.plans[g], chunk_hook, &hooks[g]);
        }
    }

    // posters feed the copies; the bands run their plans
    std::vector<std::thread> posters;
    if (resident)
        for (int g = 0; g < G; g++) posters.emplace_back(post_all, &sh, g, a->bed);
    on_each([&](int g) {
        gcta_b200_out& o = outs[g];
        memset(&o, 0, sizeof o);
        o.A_lower = out->A_lower;
        o.N_lower = out->N_lower;
        o.mu_out = g == 0 ? out->mu_out : nullptr;
        char e[512] = {0};
        // genotypes beyond the HBM allowance: every plan streams the whole file itself (no peer distribution)
        rcs[g] = gcta_b200_plan_run_to_host(sh.plans[g], resident ? nullptr : a->bed, &o, e, sizeof e);
        errs[g] = e;
        if (rcs[g]) sh.failed.store(1);
    });
    for (auto& t : posters) t.join();
    int rc = first_error();
    if (!rc && sh.failed.load()) rc = set_err(err, errcap, GCTA_B200_ERR_CUDA, "Error: make_grm_multi: a host-to-device or peer copy of the packed genotypes failed.");
    if (!rc) {
        gcta_b200_out res = outs[0];
        res.A_lower = out->A_lower; res.N_lower = out->N_lower; res.mu_out = out->mu_out;
        res.row_begin = 0;
        res.row_end = outs[G - 1].row_end;
        res.launches = 0;
        for (int g = 0; g < G; g++) {
            res.launches += outs[g].launches;
            if (outs[g].ms_total > res.ms_total) {
                res.ms_total = outs[g].ms_total; res.ms_decode = outs[g].ms_decode; res.ms_syrk = outs[g].ms_syrk;
                res.ms_paircount = outs[g].ms_paircount; res.ms_finalize = outs[g].ms_finalize;
            }
            res.total_missing = std::max(res.total_missing, outs[g].total_missing);
            if (per_gpu) per_gpu[g] = outs[g];
        }
        *out = res;
    }
    cleanup();
    return rc;
}
