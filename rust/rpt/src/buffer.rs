//! src/buffer.rs of the reference.  The Buffer is the device film (`rptx_film`): one batch mean per pixel per
//! `Renderer::sample` call, `image()` = box filter + clamp + gamma + truncating u8, `variance()` = mean per-pixel
//! unbiased variance of the batch means — computed by the library (`k_image`, `k_variance`).
use image::{ImageBuffer, RgbImage};
use rpt_cuda_sys as sys;

use crate::ffi::check;

/// A buffer that stores sample results from path tracing.  Handed to `iterative_render`'s callback by reference:
/// valid during the callback only, as in the reference.
pub struct Buffer {
    pub(crate) width: u32,
    pub(crate) height: u32,
    pub(crate) film: *mut sys::rptx_film,
    pub(crate) ctx: *mut sys::rptx_ctx,
}

impl Buffer {
    /// Converts the current buffer to an image (src/buffer.rs:43-56); panics "Pixel found with no samples"
    /// before the first batch like the reference (:89)
    pub fn image(&self) -> RgbImage {
        let mut buf = vec![0u8; (self.width * self.height * 3) as usize];
        check(self.ctx, unsafe { sys::rptx_film_image_rgb8(self.film, buf.as_mut_ptr()) });
        ImageBuffer::from_raw(self.width, self.height, buf).expect("Image buffer has incorrect size")
    }

    /// Return the average color variance of samples in each pixel (src/buffer.rs:59-73); NaN with one batch
    pub fn variance(&self) -> f64 {
        let mut v = 0.0f64;
        check(self.ctx, unsafe { sys::rptx_film_variance(self.film, &mut v) });
        v
    }

    /// Linear, unfiltered mean image [height][width][3] (addition: what the relative-RMSE checks compare)
    pub fn mean(&self) -> Vec<f64> {
        let mut out = vec![0.0f64; (self.width * self.height * 3) as usize];
        check(self.ctx, unsafe { sys::rptx_film_mean(self.film, out.as_mut_ptr()) });
        out
    }
}

impl Drop for Buffer {
    fn drop(&mut self) {
        unsafe { sys::rptx_film_destroy(self.film) }
    }
}

/// A noise reduction filter applied to the rendered image (src/buffer.rs:96-108)
#[derive(Copy, Clone)]
pub enum Filter {
    /// Box filter with a given radius
    Box(u32),
}

impl Default for Filter {
    fn default() -> Self {
        Self::Box(0)
    }
}
