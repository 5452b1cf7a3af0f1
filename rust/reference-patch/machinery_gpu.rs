//! Replaces the body of `src/renderer/machinery.rs` of bluecube/minipath: same `render()` signature and contract
//! (machinery.rs:20-30: returns immediately; callbacks fire from worker threads, `started` before a tile renders,
//! `finished` after its pixels are in the image; `abort()` lets in-flight tiles finish), work done by libminipath_b200.
//! `RenderSettings` gains `max_bounces: u32` (0 = the reference shader, worker.rs:57-64) and `seed: Option<u64>`
//! (None = OS-random like SmallRng::from_os_rng, worker.rs:25).  NOT compiled in the minipath_b200 repository.
use std::{ffi::c_void, sync::{Arc, Mutex}, time::Duration};

use image::RgbaImage;
use minipath_b200_sys as sys;

use crate::{camera::Camera, renderer::RenderSettings, scene::{triangle_bvh::{FlatBvh, TriangleBvh}, Scene}, screen_block::ScreenBlock};

#[derive(Copy, Clone, Debug)]
pub struct RenderProgressSnapshot {
    pub finished: usize,
    pub total: usize,
}

pub struct RenderProgress {
    handle: *mut sys::MpRenderProgress,
    image: Mutex<RgbaImage>,
    _scene: Arc<Scene<TriangleBvh>>,
    _flat: Arc<FlatBvh>,
    _callbacks: Box<dyn std::any::Any + Send + Sync>,
}
unsafe impl Send for RenderProgress {}
unsafe impl Sync for RenderProgress {}

struct Callbacks<F1, F2>(F1, F2);

unsafe extern "C" fn on_started<F1: Fn(ScreenBlock), F2>(user: *mut c_void, t: sys::MpTile) {
    let cb = &*(user as *const Callbacks<F1, F2>);
    (cb.0)(ScreenBlock::new([t.min_x, t.min_y].into(), [t.max_x, t.max_y].into()));
}
unsafe extern "C" fn on_finished<F1, F2: Fn(ScreenBlock, RenderProgressSnapshot)>(user: *mut c_void, t: sys::MpTile, s: sys::MpProgressSnapshot) {
    let cb = &*(user as *const Callbacks<F1, F2>);
    (cb.1)(ScreenBlock::new([t.min_x, t.min_y].into(), [t.max_x, t.max_y].into()), RenderProgressSnapshot { finished: s.finished, total: s.total });
}

fn ok(rc: i32) -> anyhow::Result<()> {
    sys::check(rc).map_err(|(code, msg)| anyhow::anyhow!("minipath_b200 ({code}): {msg}"))
}

pub fn render<F1, F2>(scene: Arc<Scene<TriangleBvh>>, camera: Camera, settings: RenderSettings, started_tile_callback: F1, finished_tile_callback: F2) -> anyhow::Result<RenderProgress>
where
    F1: Fn(ScreenBlock) + Send + Sync + 'static,
    F2: Fn(ScreenBlock, RenderProgressSnapshot) + Send + Sync + 'static,
{
    let flat = Arc::new(scene.object.flatten()); // once per scene; cache it on the Scene when rendering repeatedly
    let callbacks = Box::new(Callbacks(started_tile_callback, finished_tile_callback));
    let user = &*callbacks as *const Callbacks<F1, F2> as *mut c_void;
    let cam = camera.to_ffi(); // the 17 fields of camera.rs:9-23 → sys::MpCamera
    let st = sys::MpRenderSettings {
        tile_size: settings.tile_size.get(),
        sample_count: settings.sample_count.get(),
        resolution_x: settings.resolution.x,
        resolution_y: settings.resolution.y,
        max_bounces: settings.max_bounces,
        seed: settings.seed.unwrap_or(0),
        mode: sys::MP_MODE_AUTO, // pipeline and traversal kernel: chosen by the library from the scene and max_bounces
        ..Default::default()
    };
    let mut handle = std::ptr::null_mut();
    ok(unsafe { sys::mp_render(&flat.raw, &cam, &st, std::ptr::null(), 0, Some(on_started::<F1, F2>), Some(on_finished::<F1, F2>), user, &mut handle) })?;
    Ok(RenderProgress {
        handle,
        image: Mutex::new(RgbaImage::new(settings.resolution.x, settings.resolution.y)),
        _scene: scene,
        _flat: flat,
        _callbacks: callbacks,
    })
}

impl RenderProgress {
    pub fn progress(&self) -> RenderProgressSnapshot {
        let mut s = sys::MpProgressSnapshot::default();
        unsafe { sys::mp_progress_snapshot(self.handle, &mut s) };
        RenderProgressSnapshot { finished: s.finished, total: s.total }
    }
    pub fn is_finished(&self) -> bool {
        let mut f = 0;
        unsafe { sys::mp_progress_is_finished(self.handle, &mut f) };
        f != 0
    }
    pub fn elapsed(&self) -> Duration {
        let mut ms = 0.0f64;
        unsafe { sys::mp_progress_elapsed_ms(self.handle, &mut ms) };
        Duration::from_secs_f64(ms / 1e3)
    }
    pub fn abort(&self) {
        unsafe { sys::mp_progress_abort(self.handle) };
    }
    /// Joins the worker threads.  The reference's wait() cannot fail; a device error during the render is reported here instead of being dropped.
    pub fn wait(&self) {
        if let Err(e) = ok(unsafe { sys::mp_progress_wait(self.handle) }) {
            eprintln!("render failed: {e}");
        }
    }
    /// The image behind its mutex (machinery.rs:175-177), refreshed from the library's pinned copy under the library's image lock.
    pub fn image(&self) -> &Mutex<RgbaImage> {
        {
            let mut img = self.image.lock().unwrap();
            let cap = img.as_raw().len();
            unsafe { sys::mp_progress_image(self.handle, img.as_mut_ptr(), cap) };
        }
        &self.image
    }
}

impl Drop for RenderProgress {
    fn drop(&mut self) {
        unsafe {
            sys::mp_progress_abort(self.handle);
            sys::mp_progress_wait(self.handle);
            sys::mp_progress_free(self.handle);
        }
    }
}

// Worker (worker.rs) and StackCache are not used on this path any more; they stay for Object::intersect on single rays.
