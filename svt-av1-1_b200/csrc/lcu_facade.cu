// lcu_facade.cu -- host side of include/svtme_lcu.h: the per-SB call shape of the reference's MotionEstimateLcu
// (EbMotionEstimation.c:6152) on top of the picture-level CUDA path. No kernels here; everything it launches goes
// through svtme_picture_upload / svtme_motion_estimate (api.cu).
#include <condition_variable>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/svtme_lcu.h"

namespace {

struct Resident {
    uint64_t picture_number = 0;
    uint64_t last_use = 0;
    bool valid = false;
};

struct PictureRows {
    std::vector<SvtMePuResult> rows;     // [n_sb][num_pus]
    SvtMeParams params;
    int32_t fetched = 0;
    uint64_t born = 0;
    // The first SB call of a picture (its "leader") submits the picture and waits for the rows OUTSIDE the session lock;
    // the other callers of that picture sleep on `cv` meanwhile and other pictures can be submitted behind it.
    enum { kSubmitted, kReady, kFailed } state = kSubmitted;
    int32_t rc = SVTME_OK;
};

}  // namespace

struct SvtMeLcuSession {
    SvtMeContext *ctx = nullptr;
    int32_t width = 0, height = 0, n_sb = 0;
    std::mutex mu;
    std::vector<Resident> slots;
    std::map<uint64_t, std::shared_ptr<PictureRows>> pictures;   // keyed by picture_number of the current picture
    std::condition_variable cv;                     // a picture changed state
    uint64_t tick = 0;
    int64_t n_pictures = 0, n_uploads = 0, n_rows = 0;
    std::string err;
};

static const int kMaxPendingPictures = 64;           // rows of pictures whose SBs were never all fetched are dropped oldest-first

// Slot holding `pic` (uploading it if it is not resident); `pinned` are slots this job already uses.
static int32_t resident_slot(SvtMeLcuSession *s, const SvtMeLcuPicture *pic, const int *pinned, int n_pinned, int *slot_out) {
    int victim = -1;
    for (int i = 0; i < (int)s->slots.size(); i++) {
        Resident &r = s->slots[i];
        if (r.valid && r.picture_number == pic->picture_number) {
            r.last_use = ++s->tick;
            *slot_out = i;
            return SVTME_OK;
        }
        bool is_pinned = false;
        for (int k = 0; k < n_pinned; k++) is_pinned |= pinned[k] == i;
        if (is_pinned) continue;
        if (victim < 0) { victim = i; continue; }
        const Resident &v = s->slots[victim];            // an empty slot first, else the least recently used one
        if (v.valid && (!r.valid || r.last_use < v.last_use)) victim = i;
    }
    if (victim < 0) { s->err = "no free picture slot"; return SVTME_ERR_INSUFFICIENT_RESOURCES; }
    if (!pic->luma || pic->stride < s->width) { s->err = "bad picture plane"; return SVTME_ERR_BAD_PARAMETER; }
    const int32_t rc = svtme_picture_upload(s->ctx, victim, pic->luma, pic->stride);
    if (rc != SVTME_OK) { s->err = svtme_last_error(s->ctx); s->slots[victim].valid = false; return rc; }
    s->slots[victim].picture_number = pic->picture_number;
    s->slots[victim].last_use = ++s->tick;
    s->slots[victim].valid = true;
    s->n_uploads++;
    *slot_out = victim;
    return SVTME_OK;
}

static int32_t run_picture(SvtMeLcuSession *s, const SvtMeParams *params, const SvtMeLcuPicture *cur, const SvtMeLcuPicture *ref0,
                           const SvtMeLcuPicture *ref1, PictureRows *out) {
    int used[3], n_used = 0;
    int32_t rc;
    if ((rc = resident_slot(s, cur, used, n_used, &used[0])) != SVTME_OK) return rc;
    n_used = 1;
    if ((rc = resident_slot(s, ref0, used, n_used, &used[1])) != SVTME_OK) return rc;
    n_used = 2;
    used[2] = used[1];
    if (params->num_lists == 2 && (rc = resident_slot(s, ref1, used, n_used, &used[2])) != SVTME_OK) return rc;
    out->rows.resize((size_t)s->n_sb * params->num_pus);
    out->params = *params;
    SvtMeJob job;
    std::memset(&job, 0, sizeof(job));
    job.cur_slot = used[0]; job.ref0_slot = used[1]; job.ref1_slot = used[2];
    job.params = *params;
    job.results = out->rows.data();
    rc = svtme_motion_estimate_async(s->ctx, &job, 1);   // enqueued; the rows arrive on the context's result stream
    if (rc != SVTME_OK) s->err = svtme_last_error(s->ctx);
    return rc;
}

extern "C" {

int32_t svtme_lcu_session_create(SvtMeLcuSession **session, int32_t device, int32_t width, int32_t height, int32_t num_slots) {
    if (!session) return SVTME_ERR_BAD_PARAMETER;
    *session = nullptr;
    if (num_slots < 3) return SVTME_ERR_BAD_PARAMETER;
    SvtMeLcuSession *s = new SvtMeLcuSession;
    const int32_t rc = svtme_create(&s->ctx, device, width, height, num_slots);
    if (rc != SVTME_OK) {                           // no device / bad geometry: fail, never fall back
        if (s->ctx) svtme_destroy(s->ctx);
        delete s;
        return rc;
    }
    s->width = width; s->height = height;
    s->n_sb = ((width + 63) / 64) * ((height + 63) / 64);
    s->slots.resize(num_slots);
    *session = s;
    return SVTME_OK;
}

void svtme_lcu_session_destroy(SvtMeLcuSession *s) {
    if (!s) return;
    if (s->ctx) svtme_destroy(s->ctx);
    delete s;
}

const char *svtme_lcu_last_error(const SvtMeLcuSession *s) { return s ? s->err.c_str() : "no session"; }

int32_t svtme_lcu_motion_estimate(SvtMeLcuSession *s, const SvtMeParams *params, const SvtMeLcuPicture *cur, const SvtMeLcuPicture *ref0,
                                  const SvtMeLcuPicture *ref1, int32_t sb_index, SvtMePuResult *row_out) {
    if (!s || !params || !cur || !ref0 || !row_out) return SVTME_ERR_BAD_PARAMETER;
    std::unique_lock<std::mutex> lock(s->mu);
    if (sb_index < 0 || sb_index >= s->n_sb || params->width != s->width || params->height != s->height ||
        (params->num_lists == 2 && !ref1)) {
        s->err = "bad SB index / geometry / missing list-1 reference";
        return SVTME_ERR_BAD_PARAMETER;
    }
    std::shared_ptr<PictureRows> pr;
    auto it = s->pictures.find(cur->picture_number);
    if (it == s->pictures.end()) {
        if ((int)s->pictures.size() >= kMaxPendingPictures) {
            auto oldest = s->pictures.end();
            for (auto p = s->pictures.begin(); p != s->pictures.end(); ++p)
                if (p->second->state != PictureRows::kSubmitted && (oldest == s->pictures.end() || p->second->born < oldest->second->born)) oldest = p;
            if (oldest != s->pictures.end()) s->pictures.erase(oldest);
        }
        // leader of this picture: uploads + launch sequence are enqueued under the lock (the context is not thread-safe) ...
        pr = std::make_shared<PictureRows>();
        pr->born = ++s->tick;
        const int32_t rc = run_picture(s, params, cur, ref0, ref1, pr.get());
        if (rc != SVTME_OK) return rc;
        s->n_pictures++;
        s->pictures.emplace(cur->picture_number, pr);
        // ... and the wait for the rows is not: other ME threads submit THEIR pictures behind this one meanwhile (the GPU runs
        // them back to back), and threads asking for other SBs of this picture sleep on the condition variable
        lock.unlock();
        const int32_t wrc = svtme_fetch_results(s->ctx, nullptr, 0);
        lock.lock();
        pr->rc = wrc;
        pr->state = wrc == SVTME_OK ? PictureRows::kReady : PictureRows::kFailed;
        if (wrc != SVTME_OK) s->err = svtme_last_error(s->ctx);
        s->cv.notify_all();
    } else {
        pr = it->second;
        if (std::memcmp(&pr->params, params, sizeof(SvtMeParams)) != 0) {
            s->err = "search parameters changed between the SBs of one picture";
            return SVTME_ERR_BAD_PARAMETER;
        }
        s->cv.wait(lock, [&] { return pr->state != PictureRows::kSubmitted; });
    }
    if (pr->state == PictureRows::kFailed) return pr->rc;
    std::memcpy(row_out, pr->rows.data() + (size_t)sb_index * params->num_pus, sizeof(SvtMePuResult) * params->num_pus);
    s->n_rows++;
    if (++pr->fetched == s->n_sb) s->pictures.erase(cur->picture_number);
    return SVTME_OK;
}

/* The same for a rectangle of SBs -- one ME segment of MotionEstimationKernel (EbMotionEstimationProcess.c:432-439, 478-556):
 * SB columns [sb_x0, sb_x1) x SB rows [sb_y0, sb_y1) of the picture. rows_out[i] receives the row of the i-th SB of the rectangle
 * in raster order (the caller passes PictureParentControlSet_t::me_results[sb_index] pointers); rc_me_distortion, when not
 * NULL, receives the matching sums of the sixteen 16x16 distortions (EbMotionEstimation.c:7150-7157). One lock and one wait
 * per segment instead of one per SB. */
int32_t svtme_lcu_motion_estimate_segment(SvtMeLcuSession *s, const SvtMeParams *params, const SvtMeLcuPicture *cur, const SvtMeLcuPicture *ref0,
                                          const SvtMeLcuPicture *ref1, int32_t sb_x0, int32_t sb_x1, int32_t sb_y0, int32_t sb_y1,
                                          SvtMePuResult *const *rows_out, uint32_t *rc_me_distortion) {
    if (!s || !params || !cur || !ref0 || !rows_out) return SVTME_ERR_BAD_PARAMETER;
    const int32_t sb_cols = (s->width + 63) / 64, sb_rows = (s->height + 63) / 64;
    std::unique_lock<std::mutex> lock(s->mu);
    if (sb_x0 < 0 || sb_x1 > sb_cols || sb_x0 >= sb_x1 || sb_y0 < 0 || sb_y1 > sb_rows || sb_y0 >= sb_y1 || params->width != s->width ||
        params->height != s->height || (params->num_lists == 2 && !ref1)) {
        s->err = "bad SB rectangle / geometry / missing list-1 reference";
        return SVTME_ERR_BAD_PARAMETER;
    }
    std::shared_ptr<PictureRows> pr;
    auto it = s->pictures.find(cur->picture_number);
    if (it == s->pictures.end()) {
        if ((int)s->pictures.size() >= kMaxPendingPictures) {
            auto oldest = s->pictures.end();
            for (auto p = s->pictures.begin(); p != s->pictures.end(); ++p)
                if (p->second->state != PictureRows::kSubmitted && (oldest == s->pictures.end() || p->second->born < oldest->second->born)) oldest = p;
            if (oldest != s->pictures.end()) s->pictures.erase(oldest);
        }
        pr = std::make_shared<PictureRows>();
        pr->born = ++s->tick;
        const int32_t rc = run_picture(s, params, cur, ref0, ref1, pr.get());
        if (rc != SVTME_OK) return rc;
        s->n_pictures++;
        s->pictures.emplace(cur->picture_number, pr);
        lock.unlock();                                   // the wait is outside the lock: see svtme_lcu_motion_estimate
        const int32_t wrc = svtme_fetch_results(s->ctx, nullptr, 0);
        lock.lock();
        pr->rc = wrc;
        pr->state = wrc == SVTME_OK ? PictureRows::kReady : PictureRows::kFailed;
        if (wrc != SVTME_OK) s->err = svtme_last_error(s->ctx);
        s->cv.notify_all();
    } else {
        pr = it->second;
        if (std::memcmp(&pr->params, params, sizeof(SvtMeParams)) != 0) {
            s->err = "search parameters changed between the segments of one picture";
            return SVTME_ERR_BAD_PARAMETER;
        }
        s->cv.wait(lock, [&] { return pr->state != PictureRows::kSubmitted; });
    }
    if (pr->state == PictureRows::kFailed) return pr->rc;
    const int32_t n = (sb_x1 - sb_x0) * (sb_y1 - sb_y0);
    lock.unlock();                                       // the rows are immutable once ready; the copies need no lock
    int32_t i = 0;
    for (int32_t y = sb_y0; y < sb_y1; y++)
        for (int32_t x = sb_x0; x < sb_x1; x++, i++) {
            const SvtMePuResult *row = pr->rows.data() + (size_t)(y * sb_cols + x) * params->num_pus;
            std::memcpy(rows_out[i], row, sizeof(SvtMePuResult) * params->num_pus);
            if (rc_me_distortion) rc_me_distortion[i] = svtme_rc_me_distortion(row);
        }
    lock.lock();
    s->n_rows += n;
    pr->fetched += n;
    if (pr->fetched == s->n_sb) s->pictures.erase(cur->picture_number);
    return SVTME_OK;
}

void svtme_lcu_session_stats(const SvtMeLcuSession *s, int64_t *pictures, int64_t *uploads, int64_t *rows) {
    if (!s) return;
    if (pictures) *pictures = s->n_pictures;
    if (uploads) *uploads = s->n_uploads;
    if (rows) *rows = s->n_rows;
}

}  // extern "C"
