# B200Backend.jl -- the glue a GrayScott.jl maintainer loads to get `backend = "B200"`.
#
# It adds methods to GrayScott.Simulation from OUTSIDE the package, exactly the way
# ext/CUDAExt.jl:16,31,127,199 does for `backend = "CUDA"`; nothing in the reference is edited.
# `settings.backend = "B200"` is lower-cased to the Symbol :b200 by Inputs.load_backend_and_lang
# (src/simulation/Inputs.jl:110-120) and reaches these methods through the Val dispatch of
# src/simulation/public.jl:18-25,45-53.
#
# NOT EXECUTED in this repository's CI: the build machines have no Julia (SURVEY.md K13).  The same
# C ABI is exercised by the Python mirror grayscott.jl_b200/simulation.py, test for test.
#
#   julia --project -e 'include("julia/B200Backend.jl"); GrayScott.main(["cfg.toml"])'
#
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
Opaque device array of one `Fields` slot.  The library owns the memory (padded layout); the
wrapper only remembers which slot it names NOW: `iterate!` swaps the wrappers' slots like
public.jl:67-68 swaps the references.  `Array(a)` is the ghosted device-to-host copy.
"""
mutable struct B200Array{T, N} <: AbstractArray{T, N}
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

function Simulation.init_fields(::Val{:b200}, ::Val{:plain}, settings::Settings, mcd::MPICartDomain, T)
    rank = MPI.Comm_rank(mcd.cart_comm)
    nranks = MPI.Comm_size(mcd.cart_comm)
    ngpu = Ref{Cint}(1)   # one rank per GPU of the node, like ext/CUDAExt.jl's device selection
    cfg = GsbConfig(gsb_dtype(T), Int32(get(ENV, "GSB_ARITH", "0") == "1" ? 1 : 0), settings.L, (0, 0, 0),
                    Tuple(Int32.(mcd.dims)), Tuple(Int32.(mcd.coords)), Tuple(mcd.proc_sizes),
                    Tuple(mcd.proc_offsets),
                    Tuple(Int32(n == MPI.PROC_NULL ? -1 : n) for n in (mcd.proc_neighbors[k] for k in NEIGHBORS)),
                    settings.Du, settings.Dv, settings.F, settings.k, settings.dt, settings.noise,
                    UInt64(0x5EED), Int32(rank % parse(Int, get(ENV, "GSB_GPUS_PER_NODE", "8"))), Int32(rank), Int32(nranks))
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

# more specific than public.jl:54: skips the generic exchange! that stages GPU arrays through host
# copies and drops the result (communication.jl:180-198)
function Simulation.iterate!(::Val{:b200}, ::Val{:plain}, fields::Fields{T, 3, <:B200Array}, settings::Settings,
                             mcd::MPICartDomain) where {T}
    check(ccall((:gsb_step, libgsb), Cint, (Ptr{Cvoid}, Cint), fields.u.handle.ptr, 1))
    # the library swapped its bindings; the wrappers keep naming u/v/u_temp/v_temp, so nothing to swap here
    return
end

function Simulation.exchange!(fields::Fields{T, 3, <:B200Array}, mcd) where {T}
    check(ccall((:gsb_exchange, libgsb), Cint, (Ptr{Cvoid},), fields.u.handle.ptr))
end

function Simulation.calculate!(::Val{:b200}, ::Val{:plain}, fields::Fields{T, 3, <:B200Array}, settings::Settings,
                               mcd::MPICartDomain) where {T}
    check(ccall((:gsb_calculate, libgsb), Cint, (Ptr{Cvoid},), fields.u.handle.ptr))
end

# called by IO.write_step! (src/simulation/IO.jl:84-85)
function Simulation.get_fields(::Val{:b200}, fields::Fields{T, 3, <:B200Array}) where {T}
    sx, sy, sz = fields.u.dims .- 2
    u = Array{T, 3}(undef, sx, sy, sz)
    v = Array{T, 3}(undef, sx, sy, sz)
    check(ccall((:gsb_get_fields, libgsb), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}), fields.u.handle.ptr, u, v))
    return u, v
end

end # module
