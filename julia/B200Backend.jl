# B200Backend.jl -- the glue a GrayScott.jl maintainer loads to get `backend = "B200"`.
#
# It adds methods to GrayScott.Simulation from OUTSIDE the package, exactly the way
# ext/CUDAExt.jl:16,31,127,199 does for `backend = "CUDA"`; nothing in the reference is edited.
# `settings.backend = "B200"` is lower-cased to the Symbol :b200 by Inputs.load_backend_and_lang
# (src/simulation/Inputs.jl:110-120) and reaches these methods through the Val dispatch of
# src/simulation/public.jl:18-25,45-53.
#
# NOT EXECUTED in this repository's CI: the build machines have no Julia (SURVEY.md K13).  What is
# checked instead: tests/test_glue_layout.py parses `struct GsbConfig` below and compares it field by
# field (order, width, offset) with the C struct, builds a ctypes structure FROM this file's definition
# and replays this file's ccall sequence with it on a GPU; the Python mirror
# grayscott.jl_b200/simulation.py exercises the same C ABI test for test.
#
#   julia --project -e 'include("julia/B200Backend.jl"); GrayScott.main(["cfg.toml"])'
#
# Environment (all optional):
#   GSB200_LIB    path of libgsb200.so (default: ../grayscott.jl_b200/libgsb200.so)
#   GSB_ARITH     "reference" (default): Float32 runs use the reference's mixed Float32/Float64
#                 promotion, bit-identical to the CPU backend at noise = 0, at ~85 % of the speed;
#                 "native": Float32 FMAs, within 1e-5 of the CPU backend (the benchmark headline)
#   GSB_SEED      noise key, decimal or 0x-hex (default 0x5EED: every run repeats the same noise;
#                 "random" draws one from the clock on rank 0 and broadcasts it, like the reference's
#                 unseeded RNG)
#   GSB_DEVICE    CUDA device ordinal of this rank (default: node-local rank modulo the device count)
module B200Backend

import MPI
import GrayScott
import GrayScott: Simulation
import GrayScott.Simulation: Settings, MPICartDomain, Fields

const libgsb = get(ENV, "GSB200_LIB", joinpath(@__DIR__, "..", "grayscott.jl_b200", "libgsb200.so"))

# struct gsb_config of include/gsb200.h, field for field
struct GsbConfig
    dtype::Int32
    arith::Int32
    L::Int64
    global_::NTuple{3, Int64}
    dims::NTuple{3, Int32}
    coords::NTuple{3, Int32}
    sizes::NTuple{3, Int64}
    offsets::NTuple{3, Int64}
    neighbors::NTuple{6, Int32}
    Du::Float64
    Dv::Float64
    F::Float64
    k::Float64
    dt::Float64
    noise::Float64
    seed::UInt64
    device::Int32
    rank::Int32
    nranks::Int32
end

gsb_error() = unsafe_string(ccall((:gsb_last_error, libgsb), Cstring, ()))
check(rc::Cint) = rc == 0 ? nothing : error("B200 backend: " * gsb_error())

mutable struct Handle
    ptr::Ptr{Cvoid}
    function Handle(ptr)
        h = new(ptr)
        finalizer(h -> (h.ptr != C_NULL && ccall((:gsb_destroy, libgsb), Cint, (Ptr{Cvoid},), h.ptr); h.ptr = C_NULL), h)
        return h
    end
end

"""
Opaque device array of one `Fields` slot.  The library owns the memory (padded layout) and the
ping-pong binding: slot GSB_U always names the CURRENT u, GSB_U_TEMP the array the next calculate!
writes.  `iterate!` and `swap!` re-bind inside the library (gsb_step / gsb_swap); the wrappers are
never exchanged on the Julia side.  `Array(a)` is the ghosted device-to-host copy.
"""
struct B200Array{T, N} <: AbstractArray{T, N}
    handle::Handle
    slot::Cint                      # GSB_U, GSB_V, GSB_U_TEMP, GSB_V_TEMP
    dims::NTuple{N, Int}
end
Base.size(a::B200Array) = a.dims
Base.getindex(a::B200Array, i...) = error("B200Array is device memory; use Array(a)")
function Base.Array(a::B200Array{T, 3}) where {T}
    host = Array{T, 3}(undef, a.dims...)
    check(ccall((:gsb_get_ghosted, libgsb), Cint, (Ptr{Cvoid}, Cint, Ptr{Cvoid}), a.handle.ptr, a.slot, host))
    return host
end

gsb_dtype(::Type{Float32}) = Int32(0)
gsb_dtype(::Type{Float64}) = Int32(1)

const NEIGHBORS = ("west", "east", "down", "up", "south", "north")   # communication.jl:90-92

gsb_arith() = lowercase(get(ENV, "GSB_ARITH", "reference")) in ("native", "1") ? Int32(1) : Int32(0)

function gsb_seed(comm)
    s = get(ENV, "GSB_SEED", "0x5EED")
    seed = Ref{UInt64}(lowercase(s) == "random" ? UInt64(time_ns()) : parse(UInt64, s))
    MPI.Bcast!(seed, 0, comm)           # one key for the whole run: the noise must not depend on the rank
    return seed[]
end

function gsb_device(comm)
    haskey(ENV, "GSB_DEVICE") && return parse(Int32, ENV["GSB_DEVICE"])
    node = MPI.Comm_split_type(comm, MPI.COMM_TYPE_SHARED, MPI.Comm_rank(comm))
    n = Ref{Cint}(0)
    check(ccall((:gsb_device_count, libgsb), Cint, (Ptr{Cint},), n))
    return Int32(MPI.Comm_rank(node) % max(n[], 1))
end

function Simulation.init_fields(::Val{:b200}, ::Val{:plain}, settings::Settings, mcd::MPICartDomain, T)
    rank = MPI.Comm_rank(mcd.cart_comm)
    nranks = MPI.Comm_size(mcd.cart_comm)
    cfg = GsbConfig(gsb_dtype(T), gsb_arith(), settings.L, (0, 0, 0),
                    Tuple(Int32.(mcd.dims)), Tuple(Int32.(mcd.coords)), Tuple(mcd.proc_sizes),
                    Tuple(mcd.proc_offsets),
                    Tuple(Int32(n == MPI.PROC_NULL ? -1 : n) for n in (mcd.proc_neighbors[k] for k in NEIGHBORS)),
                    settings.Du, settings.Dv, settings.F, settings.k, settings.dt, settings.noise,
                    gsb_seed(mcd.cart_comm), gsb_device(mcd.cart_comm), Int32(rank), Int32(nranks))
    out = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:gsb_create, libgsb), Cint, (Ptr{Ptr{Cvoid}}, Ref{GsbConfig}), out, cfg))
    h = Handle(out[])
    if nranks > 1
        id = zeros(UInt8, 128)
        rank == 0 && check(ccall((:gsb_get_unique_id, libgsb), Cint, (Ptr{UInt8},), id))
        MPI.Bcast!(id, 0, mcd.cart_comm)
        check(ccall((:gsb_comm_init, libgsb), Cint, (Ptr{Cvoid}, Ptr{UInt8}), h.ptr, id))
    end
    check(ccall((:gsb_init_fields, libgsb), Cint, (Ptr{Cvoid},), h.ptr))
    dims = (mcd.proc_sizes[1] + 2, mcd.proc_sizes[2] + 2, mcd.proc_sizes[3] + 2)
    arr(slot) = B200Array{T, 3}(h, Cint(slot), dims)
    # Fields wants the three MPI face datatypes (Structs.jl:90-92); built like CUDAExt.jl:68, unused here
    xy, xz, yz = Simulation.get_MPI_faces(mcd.proc_sizes..., T)
    return Fields(arr(0), arr(1), arr(2), arr(3), xy, xz, yz)
end

handle(fields::Fields) = fields.u.handle.ptr

# The reference swaps `fields.u, fields.u_temp` by hand (public.jl:67-68).  Here the binding lives in the
# library, so the wrappers must keep their roles; a caller that exchanged them would make every later
# call address the wrong arrays.  Refuse loudly instead.
function check_roles(fields::Fields)
    (fields.u.slot == 0 && fields.v.slot == 1 && fields.u_temp.slot == 2 && fields.v_temp.slot == 3) ||
        error("B200 backend: fields.u / fields.u_temp were exchanged on the Julia side; use B200Backend.swap!(fields) " *
              "(or iterate!), which re-binds inside the library")
    return nothing
end

# more specific than public.jl:54: skips the generic exchange! that stages GPU arrays through host
# copies and drops the result (communication.jl:180-198).  gsb_step = exchange -> calculate -> swap,
# the swap being the library's own re-binding: the wrappers stay as they are.
function Simulation.iterate!(::Val{:b200}, ::Val{:plain}, fields::Fields{T, 3, <:B200Array}, settings::Settings,
                             mcd::MPICartDomain) where {T}
    check_roles(fields)
    check(ccall((:gsb_step, libgsb), Cint, (Ptr{Cvoid}, Cint), handle(fields), 1))
    return
end

# The three pieces for callers that drive them separately: exchange!, calculate!, swap!.
function Simulation.exchange!(fields::Fields{T, 3, <:B200Array}, mcd) where {T}
    check_roles(fields)
    check(ccall((:gsb_exchange, libgsb), Cint, (Ptr{Cvoid},), handle(fields)))
end

function Simulation.calculate!(::Val{:b200}, ::Val{:plain}, fields::Fields{T, 3, <:B200Array}, settings::Settings,
                               mcd::MPICartDomain) where {T}
    check_roles(fields)
    check(ccall((:gsb_calculate, libgsb), Cint, (Ptr{Cvoid},), handle(fields)))
end

"""
    swap!(fields)

The `fields.u, fields.u_temp = fields.u_temp, fields.u` of public.jl:67-68 for this backend: gsb_swap
re-binds u <-> u_temp and v <-> v_temp inside the library and advances the step (= noise) counter.
After it `fields.u` names the new values again.  Do not exchange the wrappers by hand.
"""
function swap!(fields::Fields{T, 3, <:B200Array}) where {T}
    check_roles(fields)
    check(ccall((:gsb_swap, libgsb), Cint, (Ptr{Cvoid},), handle(fields)))
end

# called by IO.write_step! (src/simulation/IO.jl:84-85): blocking, returns fresh host arrays
function Simulation.get_fields(::Val{:b200}, fields::Fields{T, 3, <:B200Array}) where {T}
    check_roles(fields)
    sx, sy, sz = fields.u.dims .- 2
    u = Array{T, 3}(undef, sx, sy, sz)
    v = Array{T, 3}(undef, sx, sy, sz)
    check(ccall((:gsb_get_fields, libgsb), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}), handle(fields), u, v))
    return u, v
end

"""
    get_fields_async!(fields, u, v); ...; output_wait(fields)

Opt-in output path for a driver that can keep stepping while the copy drains: `u`, `v` are
caller-owned host arrays of size (sx, sy, sz) that must stay alive (and should be pinned) until
`output_wait` returns.  IO.write_step! as written needs the arrays at once, so it uses get_fields.
"""
function get_fields_async!(fields::Fields{T, 3, <:B200Array}, u::Array{T, 3}, v::Array{T, 3}) where {T}
    check_roles(fields)
    check(ccall((:gsb_get_fields_async, libgsb), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}), handle(fields), u, v))
end
output_wait(fields::Fields) = check(ccall((:gsb_output_wait, libgsb), Cint, (Ptr{Cvoid},), handle(fields)))

end # module
