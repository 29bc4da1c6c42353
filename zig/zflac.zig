//! zflac.zig — drop-in replacement for Senryoku/zflac's `src/zflac.zig`: the same public surface
//!
//!     pub fn decode(allocator: std.mem.Allocator, reader: anytype) !DecodedFLAC      (reference src/zflac.zig:217)
//!     pub const DecodedFLAC = struct { channels, sample_rate, bits_per_sample, samples, _samples_backing, deinit }
//!                                                                                     (reference src/zflac.zig:12-28)
//!
//! with the body replaced by calls into libzflacb200.so (include/zflacb200.h). Everything per byte or per sample runs
//! on the GPU; there is no CPU decode fallback (a CUDA failure is `error.CudaDecodeFailed`).
//!
//! UNVERIFIED SOURCE: no Zig toolchain exists in the image this library is built and tested in. Every call made here
//! is exercised through the same C ABI by tests/test_gpu_*.py (ctypes); `zig/build.zig.fragment` shows the build wiring.
const std = @import("std");
const c = @cImport(@cInclude("zflacb200.h"));

pub const DecodedFLAC = struct { // unchanged: reference src/zflac.zig:12-28
    const Samples = union(enum) { s8: []align(32) i8, s16: []align(32) i16, s32: []align(32) i32 };
    channels: u8,
    sample_rate: u24,
    bits_per_sample: u8,
    samples: Samples,
    _samples_backing: []align(32) u8,

    pub fn deinit(self: *@This(), allocator: std.mem.Allocator) void {
        allocator.free(self._samples_backing); // "Caller owns the returned memory", reference :216, :25-27
    }
};

/// The reference's error set (sites in SURVEY.md 8b) plus the two outcomes only this backend has.
pub const Error = error{
    InvalidSignature, // :220
    InvalidMetadataHeader, // :248
    Unimplemented, // :263
    InvalidChecksum, // :280
    MissingStreaminfo, // :309
    InvalidFrameHeader, // :352,357,361,372,405
    InvalidCodedNumber, // :206
    InconsistentParameters, // :386,391
    InvalidSubframeHeader, // :431,471,542
    InvalidResidualCodingMethod, // :618
    EndOfStream,
    OutOfMemory,
    /// input on which the reference itself is undefined (unreachable, out-of-range @intCast: a panic in Debug, UB in
    /// ReleaseFast; SURVEY.md 7, Q8/Q9/Q16): reported instead of reproduced
    UndefinedInReference,
    /// no device, a kernel launch or a copy failed: zflacb200_last_error() has the detail. There is no CPU fallback.
    CudaDecodeFailed,
};

fn toError(status: c_int) Error {
    return switch (status) {
        c.ZFLACB200_INVALID_SIGNATURE => error.InvalidSignature,
        c.ZFLACB200_INVALID_METADATA_HEADER => error.InvalidMetadataHeader,
        c.ZFLACB200_UNIMPLEMENTED => error.Unimplemented,
        c.ZFLACB200_INVALID_CHECKSUM => error.InvalidChecksum,
        c.ZFLACB200_MISSING_STREAMINFO => error.MissingStreaminfo,
        c.ZFLACB200_INVALID_FRAME_HEADER => error.InvalidFrameHeader,
        c.ZFLACB200_INVALID_CODED_NUMBER => error.InvalidCodedNumber,
        c.ZFLACB200_INCONSISTENT_PARAMETERS => error.InconsistentParameters,
        c.ZFLACB200_INVALID_SUBFRAME_HEADER => error.InvalidSubframeHeader,
        c.ZFLACB200_INVALID_RESIDUAL_CODING_METHOD => error.InvalidResidualCodingMethod,
        c.ZFLACB200_END_OF_STREAM => error.EndOfStream,
        c.ZFLACB200_OUT_OF_MEMORY => error.OutOfMemory,
        c.ZFLACB200_UNDEFINED_IN_REFERENCE => error.UndefinedInReference,
        else => error.CudaDecodeFailed,
    };
}

fn result(info: c.zflacb200_info, backing: []align(32) u8) DecodedFLAC {
    const n: usize = @intCast(info.n_samples);
    return .{
        .channels = info.channels,
        .sample_rate = @intCast(info.sample_rate),
        .bits_per_sample = info.bits_per_sample,
        .samples = switch (info.container_bytes) { // union tag from STREAMINFO, reference :259-264
            1 => .{ .s8 = @as([*]align(32) i8, @ptrCast(backing.ptr))[0..n] },
            2 => .{ .s16 = @as([*]align(32) i16, @ptrCast(backing.ptr))[0..n] },
            else => .{ .s32 = @as([*]align(32) i32, @ptrCast(backing.ptr))[0..n] },
        },
        ._samples_backing = backing,
    };
}

pub fn decode(allocator: std.mem.Allocator, reader: anytype) !DecodedFLAC {
    // The GPU needs the whole stream: slurp the reader (the reference consumes it incrementally and leaves it just
    // past the last frame; this leaves it at EOF).
    const data = try reader.readAllAlloc(allocator, std.math.maxInt(usize));
    defer allocator.free(data);
    var info: c.zflacb200_info = undefined;

    // 1. STREAMINFO states a total (the normal case): size the caller-owned buffer from it and take the ONE-CALL path,
    //    which overlaps upload, kernels, read-back and the MD5 gate in chunks (13 ms instead of 18 for an hour of CD audio).
    var bound: u64 = 0;
    if (c.zflacb200_pcm_bound(data.ptr, data.len, &bound) == c.ZFLACB200_OK and bound != 0) {
        var backing = try allocator.alignedAlloc(u8, 32, @intCast(bound)); // reference :331: caller's allocator, align 32
        errdefer allocator.free(backing);
        const st = c.zflacb200_decode(data.ptr, data.len, null, backing.ptr, backing.len, &info);
        if (st == c.ZFLACB200_OK) {
            backing = try allocator.realloc(backing, @intCast(info.pcm_bytes)); // the trim of reference :583-588
            return result(info, backing);
        }
        allocator.free(backing);
        // a stream that decodes PAST its stated total (reference :394-402) does not fit: fall through to the two-phase path
        if (st != c.ZFLACB200_OUT_OF_MEMORY) return toError(st);
    }

    // 2. no stated total (or more samples than stated): decode first, then allocate exactly what came out.
    var handle: ?*c.zflacb200_stream = null;
    const st = c.zflacb200_open(data.ptr, data.len, null, &handle, &info);
    if (st != c.ZFLACB200_OK) return toError(st); // nothing is returned on error, like errdefer :265,:332
    defer c.zflacb200_close(handle);
    const backing = try allocator.alignedAlloc(u8, 32, @intCast(info.pcm_bytes));
    errdefer allocator.free(backing);
    const rd = c.zflacb200_read(handle, backing.ptr, backing.len);
    if (rd != c.ZFLACB200_OK) return toError(rd);
    return result(info, backing);
}
