//! Thin safe layer over rpt-cuda-sys: the process's context (one CUDA device), error -> panic translation.
//! Every non-zero status becomes the panic the reference raises at the same place (src/photon.rs:798,
//! src/buffer.rs:89) or a panic with the library's message; nothing aborts across the FFI.
use std::ffi::CStr;
use std::ptr;
use std::sync::{Mutex, MutexGuard};

use rpt_cuda_sys as sys;

/// rptx_ctx: one CUDA device + its stream.  Calls on one context are serialised (include/rptx.h "Threading").
pub(crate) struct Ctx {
    pub(crate) raw: *mut sys::rptx_ctx,
}

unsafe impl Send for Ctx {}

impl Drop for Ctx {
    fn drop(&mut self) {
        unsafe { sys::rptx_ctx_destroy(self.raw) }
    }
}

static CTX: Mutex<Option<Ctx>> = Mutex::new(None);

/// The process-wide context on device `RPT_CUDA_DEVICE` (default: LOCAL_RANK, else 0), created on first use.
pub(crate) fn context() -> MutexGuard<'static, Option<Ctx>> {
    let mut guard = CTX.lock().unwrap();
    if guard.is_none() {
        assert_eq!(unsafe { sys::rptx_abi_version() }, sys::RPTX_ABI_VERSION, "librptx.so ABI version mismatch");
        let device = std::env::var("RPT_CUDA_DEVICE").or_else(|_| std::env::var("LOCAL_RANK")).ok().and_then(|s| s.parse().ok()).unwrap_or(0);
        let mut raw: *mut sys::rptx_ctx = ptr::null_mut();
        let rc = unsafe { sys::rptx_ctx_create(device, &mut raw) };
        if rc != 0 {
            let msg = last_error(raw);
            unsafe { sys::rptx_ctx_destroy(raw) };
            panic!("rptx_ctx_create: {}", msg);
        }
        *guard = Some(Ctx { raw });
    }
    guard
}

pub(crate) fn last_error(ctx: *const sys::rptx_ctx) -> String {
    unsafe { CStr::from_ptr(sys::rptx_last_error(ctx)).to_string_lossy().into_owned() }
}

/// Status check: panics with the library's message (the reference panics in the same situations).
pub(crate) fn check(ctx: *const sys::rptx_ctx, rc: i32) {
    if rc != 0 {
        panic!("{}", last_error(ctx));
    }
}

/// Join this process's context into the library's NCCL communicator (one process per GPU).
pub fn comm_init(id: &[u8; sys::RPTX_COMM_ID_BYTES], world: i32, rank: i32) {
    let guard = context();
    let ctx = guard.as_ref().unwrap();
    check(ctx.raw, unsafe { sys::rptx_comm_init(ctx.raw, id.as_ptr() as *const _, world, rank) });
}

/// Rank 0 makes the communicator id and hands it to the other ranks (file, socket, MPI, ...).
pub fn comm_unique_id() -> [u8; sys::RPTX_COMM_ID_BYTES] {
    let mut id = [0u8; sys::RPTX_COMM_ID_BYTES];
    assert_eq!(unsafe { sys::rptx_comm_unique_id(id.as_mut_ptr() as *mut _) }, 0, "rptx_comm_unique_id failed (libnccl.so.2 not loadable?)");
    id
}

/// Give the device memory the context keeps cached between renders (block cache, pool, spare film slab) back to the device:
/// for a host that renders scenes of very different footprints one after another.  No counterpart in rpt.
pub fn trim_device_cache() {
    let guard = context();
    let ctx = guard.as_ref().unwrap();
    check(ctx.raw, unsafe { sys::rptx_ctx_trim(ctx.raw) });
}

/// (world, rank) of the process's context; (1, 0) without a communicator.
pub(crate) fn comm_info(ctx: &Ctx) -> (i32, i32) {
    let (mut world, mut rank) = (1, 0);
    unsafe { sys::rptx_comm_info(ctx.raw, &mut world, &mut rank) };
    (world, rank)
}
