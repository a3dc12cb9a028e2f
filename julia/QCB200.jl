# QCB200.jl -- Julia shim: an AbstractIntegrator that forwards to libqcb200.so via ccall.
#
# NOT EXECUTED in this repository's CI (no Julia toolchain in the build image).  It shows the
# reference-side binding a maintainer adds; the method names that DirectTrajOpt 0.8 dispatches on
# for Jacobian / Hessian values must be checked against the installed package (SURVEY.md 8b marks
# them [memory]); only the ccall signatures below are normative (include/qcb200.h).
module QCB200

using DirectTrajOpt
using NamedTrajectories
using PiccoloQuantumObjects
import DirectTrajOpt: evaluate!

const LIB = get(ENV, "QCB200_LIB", "libqcb200.so")

struct QcbDesc                      # mirrors qcb_desc, include/qcb200.h
    struct_bytes::Int32; d::Int32; c::Int32; m::Int32; N::Int32; S::Int32; z_dim::Int32
    u_off::Int32; dt_off::Int32; jac_mode::Int32; device::Int32; k_begin::Int32; k_end::Int32
    force_path::Int32
    x_off::Ptr{Int32}; G_drift::Ptr{Float64}; G_drives::Ptr{Float64}
    x_index::Ptr{Int32}; row_index::Ptr{Int32}      # optional state index map (variational integrators)
end

mutable struct B200BilinearIntegrator <: DirectTrajOpt.AbstractIntegrator
    handle::Ptr{Cvoid}
    dim::Int                        # x_dim * (N-1) [* S]: same meaning as BilinearIntegrator.dim
    jac_nnz::Int
    hess_nnz::Int
end

check(h, st) = st == 0 || error(unsafe_string(ccall((:qcb_last_error, LIB), Cstring, (Ptr{Cvoid},), h)))

"""
    B200BilinearIntegrator(qtraj, traj; device=0)

Drop-in for `BilinearIntegrator(qtraj, N)` (src/quantum_integrators.jl:28-134).  Pass it through the
templates' own seam:  `SmoothPulseProblem(qtraj, N; integrator = B200BilinearIntegrator(qtraj, traj))`
(src/problem_templates/smooth_pulse_problem.jl:57,103-127); for `SamplingProblem`, which hard-codes
its integrators (sampling_problem.jl:153), replace `prob.integrators[i]` after construction.
"""
function B200BilinearIntegrator(systems::Vector, state_names::Vector{Symbol}, drive::Symbol,
                                traj::NamedTrajectory, columns::Int; device::Int = 0, jac_mode::Int = 0,
                                generator = (s, u) -> s.G(u, 0.0))
    m = systems[1].n_drives
    d = size(generator(systems[1], zeros(m)), 1)     # 2n (ket / unitary) or 2n^2 (density, sys.𝒢)
    # matrices are dense column-major Float64: exactly what the C ABI wants
    G0 = reduce(hcat, [vec(Matrix{Float64}(generator(s, zeros(m)))) for s in systems])
    Gd = reduce(hcat, [vec(Matrix{Float64}(generator(s, Float64.(1:m .== j))) .- reshape(G0[:, i], d, d))
                       for (i, s) in enumerate(systems) for j in 1:m])
    return create_handle(d, columns, m, traj, length(systems), state_names, drive, G0, Gd;
                         device = device, jac_mode = jac_mode)
end

function create_handle(d, columns, m, traj, S, state_names, drive, G0, Gd; device = 0, jac_mode = 0,
                       x_index::Vector{Int32} = Int32[], row_index::Vector{Int32} = Int32[])
    x_off = Int32[first(traj.components[n]) - 1 for n in state_names]
    desc = QcbDesc(sizeof(QcbDesc), d, columns, m, traj.N, S, traj.dim,
                   first(traj.components[drive]) - 1, first(traj.components[traj.timestep]) - 1,
                   jac_mode, device, 0, 0, 0, pointer(x_off), pointer(G0), pointer(Gd),
                   isempty(x_index) ? C_NULL : pointer(x_index), isempty(row_index) ? C_NULL : pointer(row_index))
    h = Ref{Ptr{Cvoid}}(C_NULL)
    GC.@preserve x_off G0 Gd x_index row_index begin
        st = ccall((:qcb_create, LIB), Cint, (Ref{Ptr{Cvoid}}, Ref{QcbDesc}), h, desc)
    end
    st == 0 || error(unsafe_string(ccall((:qcb_last_error, LIB), Cstring, (Ptr{Cvoid},), C_NULL)))
    int = B200BilinearIntegrator(h[],
        ccall((:qcb_defect_dim, LIB), Int64, (Ptr{Cvoid},), h[]),
        ccall((:qcb_jacobian_nnz, LIB), Int64, (Ptr{Cvoid},), h[]),
        ccall((:qcb_hessian_nnz, LIB), Int64, (Ptr{Cvoid},), h[]))
    finalizer(i -> ccall((:qcb_destroy, LIB), Cvoid, (Ptr{Cvoid},), i.handle), int)
    return int
end

# DirectTrajOpt.evaluate!(delta, integrator, traj)  (smooth_pulse_problem.jl:512-515)
function evaluate!(δ::AbstractVector{Float64}, int::B200BilinearIntegrator, traj::NamedTrajectory)
    Z = traj.datavec
    GC.@preserve δ Z check(int.handle, ccall((:qcb_eval_defect, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}), int.handle, Z, δ))
    return nothing
end

# Adopt the structure DirectTrajOpt itself reports for the stock integrator, so that values land
# in exactly the reference's COO order (bit-exact sparsity).
function adopt_structure!(int::B200BilinearIntegrator, jac_rows::Vector{Int}, jac_cols::Vector{Int},
                          hess_rows::Vector{Int}, hess_cols::Vector{Int}; row_base::Int = 0)
    GC.@preserve jac_rows jac_cols hess_rows hess_cols check(int.handle, ccall((:qcb_set_structure, LIB), Cint,
        (Ptr{Cvoid}, Int64, Ptr{Int64}, Ptr{Int64}, Int64, Ptr{Int64}, Ptr{Int64}, Int64),
        int.handle, row_base, jac_rows, jac_cols, length(jac_rows), hess_rows, hess_cols, length(hess_rows)))
    int.jac_nnz = ccall((:qcb_jacobian_nnz, LIB), Int64, (Ptr{Cvoid},), int.handle)
    int.hess_nnz = ccall((:qcb_hessian_nnz, LIB), Int64, (Ptr{Cvoid},), int.handle)
end

# Page-lock the value arrays Ipopt hands to the callbacks (same arrays every iteration): the library then DMAs
# the u / dt blocks straight into them.  Call once after the problem is built; unpin before the arrays die.
pin!(a::Vector{Float64}) = (GC.@preserve a ccall((:qcb_host_register, LIB), Cint, (Ptr{Cvoid}, Csize_t), a, sizeof(a)); a)
unpin!(a::Vector{Float64}) = (GC.@preserve a ccall((:qcb_host_unregister, LIB), Cvoid, (Ptr{Cvoid},), a); a)

# eval_jac_g / eval_h value callbacks: write straight into Ipopt's value buffers
function jacobian_values!(vals::AbstractVector{Float64}, int::B200BilinearIntegrator, Z::Vector{Float64})
    GC.@preserve vals Z check(int.handle, ccall((:qcb_eval_jacobian, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}), int.handle, Z, vals))
end
function hessian_of_lagrangian_values!(vals::AbstractVector{Float64}, int::B200BilinearIntegrator,
                                       Z::Vector{Float64}, μ::Vector{Float64})
    GC.@preserve vals Z μ check(int.handle, ccall((:qcb_eval_hessian, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}), int.handle, Z, μ, vals))
end

# The same callbacks carrying Ipopt's new_x flag (MOI passes it through eval_constraint / eval_constraint_jacobian /
# eval_hessian_lagrangian as "x changed since the last call" [memory]): with new_x = false the library reuses what
# the previous callback computed for this iterate.
function evaluate!(δ::AbstractVector{Float64}, int::B200BilinearIntegrator, Z::Vector{Float64}, new_x::Bool)
    GC.@preserve δ Z check(int.handle, ccall((:qcb_eval_defect_x, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}), int.handle, Z, new_x, δ))
end
function jacobian_values!(vals::AbstractVector{Float64}, int::B200BilinearIntegrator, Z::Vector{Float64}, new_x::Bool)
    GC.@preserve vals Z check(int.handle, ccall((:qcb_eval_jacobian_x, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}), int.handle, Z, new_x, vals))
end
function hessian_of_lagrangian_values!(vals::AbstractVector{Float64}, int::B200BilinearIntegrator,
                                       Z::Vector{Float64}, μ::Vector{Float64}, new_x::Bool)
    GC.@preserve vals Z μ check(int.handle, ccall((:qcb_eval_hessian_x, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}, Ptr{Float64}), int.handle, Z, new_x, μ, vals))
end

# Structure queries: what the MOI evaluator of DirectTrajOpt asks of an integrator before the solve.
# The generic names below are [memory] (SURVEY.md 8b); the arrays are 1-based, rows relative to
# `row_base` (the offset of this integrator's defect rows in the constraint vector).
function jacobian_structure(int::B200BilinearIntegrator; row_base::Int = 0)
    rows, cols = Vector{Int64}(undef, int.jac_nnz), Vector{Int64}(undef, int.jac_nnz)
    GC.@preserve rows cols check(int.handle, ccall((:qcb_jacobian_structure, LIB), Cint,
        (Ptr{Cvoid}, Int64, Ptr{Int64}, Ptr{Int64}), int.handle, row_base, rows, cols))
    return collect(zip(rows, cols))
end
function hessian_structure(int::B200BilinearIntegrator)
    rows, cols = Vector{Int64}(undef, int.hess_nnz), Vector{Int64}(undef, int.hess_nnz)
    GC.@preserve rows cols check(int.handle, ccall((:qcb_hessian_structure, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Int64}, Ptr{Int64}), int.handle, rows, cols))
    return collect(zip(rows, cols))
end
# [memory] names DirectTrajOpt's evaluator may dispatch on instead of the two above
const jacobian! = jacobian_values!
const hessian_of_lagrangian! = hessian_of_lagrangian_values!

# all three from one pass (one H2D of Z, one evaluation): for evaluators that cache per iterate
function evaluate_all!(δ::Vector{Float64}, jac::Vector{Float64}, hess::Vector{Float64},
                       int::B200BilinearIntegrator, Z::Vector{Float64}, μ::Vector{Float64})
    GC.@preserve δ jac hess Z μ check(int.handle, ccall((:qcb_eval_all, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
        int.handle, Z, μ, δ, jac, hess))
end

# Dispatch mirroring src/quantum_integrators.jl:28-101
B200BilinearIntegrator(q::UnitaryTrajectory, traj; kw...) =
    B200BilinearIntegrator([get_system(q)], [state_name(q)], drive_name(q), traj, get_system(q).levels; kw...)
B200BilinearIntegrator(q::KetTrajectory, traj; kw...) =
    B200BilinearIntegrator([get_system(q)], [state_name(q)], drive_name(q), traj, 1; kw...)
# DensityTrajectory (:52-56): the Lindbladian sys.𝒢 on the density iso-vec, d = 2n^2, one column
B200BilinearIntegrator(q::DensityTrajectory, traj; kw...) =
    B200BilinearIntegrator([get_system(q)], [state_name(q)], drive_name(q), traj, 1;
                           generator = (s, u) -> s.𝒢(u, 0.0), kw...)
# MultiKetTrajectory (:63-74) returns a Vector, one integrator per ket, like the reference ...
B200BilinearIntegrator(q::MultiKetTrajectory, traj; kw...) =
    [B200BilinearIntegrator([get_system(q)], [n], drive_name(q), traj, 1; kw...) for n in state_names(q)]
# ... and, fused, one handle that evaluates every ket of a knot in one launch
B200FusedIntegrator(q::MultiKetTrajectory, traj; kw...) =
    B200BilinearIntegrator(fill(get_system(q), length(state_names(q))), state_names(q), drive_name(q), traj, 1; kw...)
B200BilinearIntegrator(q::SamplingTrajectory, traj; kw...) =
    B200BilinearIntegrator(q.systems, state_names(q), drive_name(q), traj,
                           q.base_trajectory isa UnitaryTrajectory ? q.systems[1].levels : 1; kw...)

# SamplingProblem hard-codes `BilinearIntegrator(sampling_qtraj, N)` (sampling_problem.jl:153); MinimumTimeProblem
# re-uses the vector (minimum_time_problem.jl:131).  Swap the stock bilinear integrators of a built
# problem for one fused handle (all systems of a knot in one launch); other integrators stay.
function replace_integrators!(qcp, qtraj; kw...)
    ints = qcp.prob.integrators
    keep = [i for i in ints if !(i isa DirectTrajOpt.BilinearIntegrator)]
    empty!(ints)
    push!(ints, B200BilinearIntegrator(qtraj, qcp.prob.trajectory; kw...))
    append!(ints, keep)
    return qcp
end

# Variational integrators (src/quantum_integrators.jl:140-169): x = vcat(state, variations...) under
# var_G(G, G_vars / scale).  The panel stacks the blocks row-wise; x_index / row_index tell the library
# where each panel entry lives in the knot and among the defect rows (include/qcb200.h, qcb_desc).
function B200VariationalIntegrator(sys, traj::NamedTrajectory, names::Vector{Symbol}, drive::Symbol,
                                   columns::Int, scales::Vector{Float64}; device::Int = 0, jac_mode::Int = 0)
    m, d0, nb = sys.n_drives, 2 * sys.levels, length(names)
    gen(u) = Matrix{Float64}(Isomorphisms.var_G(sys.G(u), [G(u) / sc for (sc, G) in zip(scales, sys.G_vars)]))
    d = nb * d0
    G0 = vec(gen(zeros(m)))
    Gd = reduce(hcat, [vec(gen(Float64.(1:m .== j))) .- G0 for j in 1:m])
    x_index, row_index = zeros(Int32, d * columns), zeros(Int32, d * columns)
    for (b, nm) in enumerate(names), q in 1:columns, r in 1:d0
        e = (b - 1) * d0 + r + (q - 1) * d                       # panel entry, 1-based
        x_index[e] = traj.components[nm][(q - 1) * d0 + r] - 1
        row_index[e] = (b - 1) * d0 * columns + (q - 1) * d0 + r - 1
    end
    return create_handle(d, columns, m, traj, 1, [names[1]], drive, G0, Gd; device = device, jac_mode = jac_mode,
                         x_index = x_index, row_index = row_index)
end
VariationalKetIntegrator(sys, traj, ψ̃::Symbol, vars::AbstractVector{Symbol}, u::Symbol; scale::Float64 = 1.0, kw...) =
    B200VariationalIntegrator(sys, traj, vcat(ψ̃, vars...), u, 1, fill(scale, length(vars)); kw...)
VariationalUnitaryIntegrator(sys, traj, Ũ⃗::Symbol, vars::AbstractVector{Symbol}, u::Symbol;
                             scales::AbstractVector{<:Float64} = fill(1.0, length(vars)), kw...) =
    B200VariationalIntegrator(sys, traj, vcat(Ũ⃗, vars...), u, sys.levels, collect(scales); kw...)

# Terminal fidelities (include/qcb200.h: qcb_fidelity_*): the losses of src/quantum_objectives.jl
# (unitary :173-191, ket :21-27, coherent ket :88-103) as quadratic forms evaluated on the GPU.  The
# objective Q*|1-F| (:199) and the constraint F_min - F (src/quantum_constraints.jl:112) follow by the
# chain rule below, exactly as TerminalObjective / NonlinearKnotPointConstraint wrap the loss closures.
const FID_UNITARY, FID_KET, FID_COHERENT_KET, FID_UNITARY_SUBSPACE = Int32(0), Int32(1), Int32(2), Int32(3)
const FID_DENSITY_PURE = Int32(4)    # density_matrix_pure_state_infidelity_loss, src/quantum_objectives.jl:232-239

mutable struct B200Fidelity
    handle::Ptr{Cvoid}
    x_dim::Int
end

function B200Fidelity(kind::Int32, goal::AbstractArray{ComplexF64}, n::Int; count::Int=1,
                      subspace::Union{Nothing,Vector{Int}}=nothing, device::Int=0)
    g = reinterpret(Float64, vec(Matrix{ComplexF64}(reshape(goal, :, 1))))   # interleaved (re, im), column-major
    sub = isnothing(subspace) ? Int32[] : Int32.(subspace .- 1)
    h = Ref{Ptr{Cvoid}}(C_NULL)
    st = GC.@preserve g sub ccall((:qcb_fidelity_create, LIB), Cint,
        (Ref{Ptr{Cvoid}}, Int32, Int32, Int32, Ptr{Float64}, Ptr{Int32}, Int32, Int32),
        h, kind, n, count, g, isempty(sub) ? C_NULL : pointer(sub), length(sub), device)
    st == 0 || error(unsafe_string(ccall((:qcb_last_error, LIB), Cstring, (Ptr{Cvoid},), C_NULL)))
    f = B200Fidelity(h[], ccall((:qcb_fidelity_x_dim, LIB), Int64, (Ptr{Cvoid},), h[]))
    finalizer(o -> ccall((:qcb_fidelity_destroy, LIB), Cvoid, (Ptr{Cvoid},), o.handle), f)
    return f
end

# F(x) and, optionally, its gradient (x = the terminal iso-vec state, e.g. traj[end][name])
function fidelity!(grad::Union{Nothing,Vector{Float64}}, f::B200Fidelity, x::Vector{Float64})
    F = Ref{Float64}(0.0)
    GC.@preserve grad x st = ccall((:qcb_fidelity_eval, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Ref{Float64}, Ptr{Float64}, Ptr{Float64}),
        f.handle, x, F, isnothing(grad) ? C_NULL : pointer(grad), C_NULL)
    st == 0 || error("qcb_fidelity_eval failed with status $st")
    return F[]
end

# Q * |1 - F| and its gradient: what UnitaryInfidelityObjective / KetInfidelityObjective hand to Ipopt
function infidelity_objective!(grad::Vector{Float64}, f::B200Fidelity, x::Vector{Float64}; Q::Float64=100.0)
    F = fidelity!(grad, f, x)
    grad .*= -Q * sign(1 - F)
    return Q * abs(1 - F)
end

# F_min - F <= 0 and its Jacobian row: Final*FidelityConstraint
function fidelity_constraint!(jac_row::Vector{Float64}, f::B200Fidelity, x::Vector{Float64}, F_min::Float64)
    F = fidelity!(jac_row, f, x)
    jac_row .*= -1.0
    return F_min - F
end

# Objective terms around the integrator (include/qcb200.h: qcb_aux_*): quadratic regularisers and the
# minimum-time term in the form the reference's own solver logs pin (DESIGN.md section 6).
struct QcbAuxDesc
    struct_bytes::Int32; N::Int32; z_dim::Int32; dt_off::Int32; device::Int32
    n_deriv::Int32; leak_x_off::Int32; n_leak::Int32
    deriv_x_off::Ptr{Int32}; deriv_dx_off::Ptr{Int32}; deriv_dim::Ptr{Int32}; leak_idx::Ptr{Int32}
end

mutable struct B200ObjectiveTerms
    handle::Ptr{Cvoid}
    off::Vector{Int32}; dim::Vector{Int32}; R::Vector{Float64}; D::Float64
    grad_cols::Vector{Int64}      # 1-based columns of Z of the compact gradient values
    grad_vals::Vector{Float64}
end

function B200ObjectiveTerms(traj::NamedTrajectory, names::Vector{Symbol}, Rs::Vector{Float64}; D::Float64=0.0, device::Int=0)
    desc = QcbAuxDesc(Int32(sizeof(QcbAuxDesc)), traj.N, traj.dim, traj.components[traj.timestep][1] - 1, device,
                      0, 0, 0, C_NULL, C_NULL, C_NULL, C_NULL)
    h = Ref{Ptr{Cvoid}}(C_NULL)
    st = ccall((:qcb_aux_create, LIB), Cint, (Ref{Ptr{Cvoid}}, Ref{QcbAuxDesc}), h, desc)
    st == 0 || error(unsafe_string(ccall((:qcb_last_error, LIB), Cstring, (Ptr{Cvoid},), C_NULL)))
    off = Int32[traj.components[n][1] - 1 for n in names]
    dim = Int32[length(traj.components[n]) for n in names]
    R = vcat([fill(Rs[i], dim[i]) for i in eachindex(names)]...)
    ng = traj.N * (sum(dim) + 1)
    cols = Vector{Int64}(undef, ng)
    GC.@preserve off dim cols ccall((:qcb_aux_regularizer_structure, LIB), Cint,
        (Ptr{Cvoid}, Int32, Ptr{Int32}, Ptr{Int32}, Ptr{Int64}, Ptr{Int64}, Ptr{Int64}),
        h[], length(dim), off, dim, cols, C_NULL, C_NULL)
    obj = B200ObjectiveTerms(h[], off, dim, R, D, cols, zeros(ng))
    finalizer(o -> ccall((:qcb_aux_destroy, LIB), Cvoid, (Ptr{Cvoid},), o.handle), obj)
    return obj
end

# value; the gradient is ADDED into `grad` (length of Z) at the regularised entries and the timesteps --
# what Ipopt's eval_f / eval_grad_f need on top of the terminal objective
function objective_and_gradient!(grad::Vector{Float64}, o::B200ObjectiveTerms, Z::Vector{Float64})
    J = Ref{Float64}(0.0)
    GC.@preserve Z o st = ccall((:qcb_aux_regularizer_eval, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Int32}, Ptr{Int32}, Ptr{Float64}, Float64, Ref{Float64}, Ptr{Float64}, Ptr{Float64}),
        o.handle, Z, length(o.dim), o.off, o.dim, o.R, o.D, J, o.grad_vals, C_NULL)
    st == 0 || error("qcb_aux_regularizer_eval failed with status $st")
    @inbounds for i in eachindex(o.grad_cols)
        grad[o.grad_cols[i]] += o.grad_vals[i]
    end
    return J[]
end

# Free-phase fidelity (src/quantum_objectives.jl:203-226; src/quantum_constraints.jl:123-144) for
# U_goal(θ) = diag(exp(i P θ)) U_goal (side 0) or U_goal diag(...) (side 1); z = [x_N; θ]
mutable struct B200FreePhase
    handle::Ptr{Cvoid}
    x_dim::Int
    n_theta::Int
end
function B200FreePhase(U_goal::AbstractMatrix{ComplexF64}, phase_map::AbstractMatrix{Float64}; side::Int = 0, device::Int = 0)
    n = size(U_goal, 1)
    g = reinterpret(Float64, vec(Matrix{ComplexF64}(U_goal)))
    P = Matrix{Float64}(phase_map)                         # n x n_theta, column-major
    h = Ref{Ptr{Cvoid}}(C_NULL)
    st = GC.@preserve g P ccall((:qcb_free_phase_create, LIB), Cint,
        (Ref{Ptr{Cvoid}}, Int32, Ptr{Float64}, Int32, Ptr{Float64}, Int32, Int32),
        h, n, g, size(P, 2), P, side, device)
    st == 0 || error(unsafe_string(ccall((:qcb_last_error, LIB), Cstring, (Ptr{Cvoid},), C_NULL)))
    f = B200FreePhase(h[], 2 * n * n, size(P, 2))
    finalizer(o -> ccall((:qcb_free_phase_destroy, LIB), Cvoid, (Ptr{Cvoid},), o.handle), f)
    return f
end
# F and its gradient over z = [x; θ] (grad has x_dim + n_theta entries; hess, if given, (x_dim+n_theta)^2)
function free_phase_fidelity!(grad::Union{Nothing,Vector{Float64}}, f::B200FreePhase, x::Vector{Float64}, θ::Vector{Float64};
                              hess::Union{Nothing,Matrix{Float64}} = nothing)
    F = Ref{Float64}(0.0)
    GC.@preserve grad x θ hess st = ccall((:qcb_free_phase_eval, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ref{Float64}, Ptr{Float64}, Ptr{Float64}),
        f.handle, x, θ, F, isnothing(grad) ? C_NULL : pointer(grad), isnothing(hess) ? C_NULL : pointer(hess))
    st == 0 || error("qcb_free_phase_eval failed with status $st")
    return F[]
end

# UnitarySensitivityObjective (src/quantum_objectives.jl:263-279) through an aux handle `a` (Ptr from qcb_aux_create)
function sensitivity_objective!(grad_vals::Matrix{Float64}, a::Ptr{Cvoid}, Z::Vector{Float64}, x_off::Int, x_dim::Int,
                                n::Int, times::Vector{Int32}, Qs::Vector{Float64}, scale::Float64)
    J = Ref{Float64}(0.0)
    GC.@preserve grad_vals Z times Qs st = ccall((:qcb_aux_sensitivity_eval, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Int32, Int32, Int32, Int32, Ptr{Int32}, Ptr{Float64}, Float64, Ref{Float64}, Ptr{Float64}, Ptr{Float64}),
        a, Z, x_off, x_dim, n, length(times), times, Qs, scale, J, grad_vals, C_NULL)
    st == 0 || error("qcb_aux_sensitivity_eval failed with status $st")
    return J[]
end

# ---- one ccall per Ipopt callback (include/qcb200.h: qcb_problem_*) -------------------------------------------
# The whole list a problem template hands to DirectTrajOpt (smooth_pulse_problem.jl:139-165: terminal infidelity +
# regularisers; BilinearIntegrator + two DerivativeIntegrators; the timesteps-all-equal rows of
# _problem_templates.jl:38-67) behind five calls that write Ipopt's full-length arrays.  `new_x` is Ipopt's flag.
struct QcbProblemDesc               # mirrors qcb_problem_desc
    struct_bytes::Int32; N::Int32; z_dim::Int32; n_integrators::Int32
    use_derivative_rows::Int32; use_dt_equal_rows::Int32; fid_hessian::Int32; n_reg::Int32
    use_leakage_rows::Int32; use_leakage_objective::Int32; n_fid_states::Int32; reserved0::Int32
    Q::Float64; min_time_D::Float64; leak_Q::Float64; leak_value::Float64
    integrators::Ptr{Ptr{Cvoid}}; aux::Ptr{Cvoid}; fidelity::Ptr{Cvoid}
    fid_index::Ptr{Int32}; fid_weights::Ptr{Float64}; reg_off::Ptr{Int32}; reg_dim::Ptr{Int32}; reg_R::Ptr{Float64}
end

mutable struct B200Problem
    handle::Ptr{Cvoid}
    n::Int; m::Int; jac_nnz::Int; hess_nnz::Int
    keep::Any                       # the borrowed pieces and index arrays: alive as long as the problem
end

"""
    B200Problem(traj, integrators; aux, derivative_rows, dt_equal_rows, fidelity, state_name, Q, fid_hessian,
                reg_names, reg_R, D)

`integrators`: `B200BilinearIntegrator`s over the whole trajectory; `aux`: a `Ptr` from `qcb_aux_create` built with the
derivative pairs (u, du), (du, ddu); `fidelity`: a `B200Fidelity` on the component `state_name` of the last knot.
"""
function B200Problem(traj::NamedTrajectory, integrators::Vector{B200BilinearIntegrator}; aux::Ptr{Cvoid} = C_NULL,
                     derivative_rows::Bool = false, dt_equal_rows::Bool = false,
                     fidelity::Union{Nothing,B200Fidelity} = nothing, state_name::Union{Nothing,Symbol} = nothing,
                     state_names::Vector{Symbol} = isnothing(state_name) ? Symbol[] : [state_name],
                     fid_weights::Vector{Float64} = Float64[],
                     Q::Float64 = 100.0, fid_hessian::Bool = true, reg_names::Vector{Symbol} = Symbol[],
                     reg_R::Vector{Float64} = Float64[], D::Float64 = 0.0, leakage_rows::Bool = false,
                     leakage_value::Float64 = 0.0, leakage_Q::Union{Nothing,Float64} = nothing)
    hs = Ptr{Cvoid}[i.handle for i in integrators]
    fidx = isempty(state_names) ? Int32[] : Int32.(vcat([traj.components[n] for n in state_names]...) .- 1)   # one state per system
    off = Int32[traj.components[n][1] - 1 for n in reg_names]
    dim = Int32[length(traj.components[n]) for n in reg_names]
    R = isempty(reg_names) ? Float64[] : vcat([fill(reg_R[i], dim[i]) for i in eachindex(reg_names)]...)
    h = Ref{Ptr{Cvoid}}(C_NULL)
    GC.@preserve hs fidx fid_weights off dim R begin
        desc = QcbProblemDesc(Int32(sizeof(QcbProblemDesc)), traj.N, traj.dim, length(hs), derivative_rows, dt_equal_rows,
                              fid_hessian, length(dim), leakage_rows, !isnothing(leakage_Q), length(state_names), 0, Q, D,
                              isnothing(leakage_Q) ? 0.0 : leakage_Q, leakage_value, pointer(hs), aux,
                              isnothing(fidelity) ? C_NULL : fidelity.handle,
                              isempty(fidx) ? C_NULL : pointer(fidx), isempty(fid_weights) ? C_NULL : pointer(fid_weights),
                              isempty(off) ? C_NULL : pointer(off),
                              isempty(dim) ? C_NULL : pointer(dim), isempty(R) ? C_NULL : pointer(R))
        st = ccall((:qcb_problem_create, LIB), Cint, (Ref{Ptr{Cvoid}}, Ref{QcbProblemDesc}), h, desc)
        st == 0 || error(unsafe_string(ccall((:qcb_problem_last_error, LIB), Cstring, (Ptr{Cvoid},), C_NULL)))
    end
    d = [Ref{Int64}(0) for _ in 1:4]
    ccall((:qcb_problem_dims, LIB), Cint, (Ptr{Cvoid}, Ref{Int64}, Ref{Int64}, Ref{Int64}, Ref{Int64}), h[], d[1], d[2], d[3], d[4])
    prob = B200Problem(h[], d[1][], d[2][], d[3][], d[4][], (integrators, fidelity, hs, fidx, fid_weights, off, dim, R))
    finalizer(o -> ccall((:qcb_problem_destroy, LIB), Cvoid, (Ptr{Cvoid},), o.handle), prob)
    return prob
end

pcheck(p::B200Problem, st) = st == 0 || error(unsafe_string(ccall((:qcb_problem_last_error, LIB), Cstring, (Ptr{Cvoid},), p.handle)))

# the (row, column) lists MOI.jacobian_structure / MOI.hessian_lagrangian_structure return (1-based)
function jacobian_structure(p::B200Problem)
    rows = Vector{Int64}(undef, p.jac_nnz); cols = similar(rows)
    pcheck(p, ccall((:qcb_problem_jacobian_structure, LIB), Cint, (Ptr{Cvoid}, Ptr{Int64}, Ptr{Int64}), p.handle, rows, cols))
    return collect(zip(rows, cols))
end
function hessian_lagrangian_structure(p::B200Problem)
    rows = Vector{Int64}(undef, p.hess_nnz); cols = similar(rows)
    pcheck(p, ccall((:qcb_problem_hessian_structure, LIB), Cint, (Ptr{Cvoid}, Ptr{Int64}, Ptr{Int64}), p.handle, rows, cols))
    return collect(zip(rows, cols))
end
# Ipopt's five callbacks (Ipopt.jl: eval_f(x), eval_grad_f(x, grad), eval_g(x, g), eval_jac_g(x, rows, cols, values),
# eval_h(x, rows, cols, obj_factor, lambda, values)); MOI's AbstractNLPEvaluator methods have the same content
function eval_f(p::B200Problem, Z::Vector{Float64}; new_x::Bool = true)
    f = Ref{Float64}(0.0)
    pcheck(p, ccall((:qcb_problem_eval_f, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Int32, Ref{Float64}), p.handle, Z, new_x, f))
    return f[]
end
function eval_grad_f!(grad::Vector{Float64}, p::B200Problem, Z::Vector{Float64}; new_x::Bool = false)
    pcheck(p, ccall((:qcb_problem_eval_grad_f, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}), p.handle, Z, new_x, grad))
    return grad
end
function eval_g!(g::Vector{Float64}, p::B200Problem, Z::Vector{Float64}; new_x::Bool = false)
    pcheck(p, ccall((:qcb_problem_eval_g, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}), p.handle, Z, new_x, g))
    return g
end
function eval_jac_g!(vals::Vector{Float64}, p::B200Problem, Z::Vector{Float64}; new_x::Bool = false)
    pcheck(p, ccall((:qcb_problem_eval_jac_g, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}), p.handle, Z, new_x, vals))
    return vals
end
function eval_h!(vals::Vector{Float64}, p::B200Problem, Z::Vector{Float64}, obj_factor::Float64, lambda::Vector{Float64};
                 new_x::Bool = false)
    pcheck(p, ccall((:qcb_problem_eval_h, LIB), Cint, (Ptr{Cvoid}, Ptr{Float64}, Int32, Float64, Ptr{Float64}, Ptr{Float64}),
                    p.handle, Z, new_x, obj_factor, lambda, vals))
    return vals
end

# Old (docs / BASELINE.json) names, SURVEY.md Appendix A
const UnitaryIntegrator = B200BilinearIntegrator
const KetIntegrator = B200BilinearIntegrator
const DensityOperatorIntegrator = B200BilinearIntegrator
# problem templates under their docs-era names: build the trajectory object, pass this integrator through
# the template's own `integrator=` seam (smooth_pulse_problem.jl:57,103-127)
function UnitarySmoothPulseProblem(sys, U_goal, T, Δt; device::Int = 0, kwargs...)
    qtraj = UnitaryTrajectory(sys, U_goal, T * Δt)
    traj = NamedTrajectory(qtraj, T)
    return SmoothPulseProblem(qtraj, T; integrator = B200BilinearIntegrator(qtraj, traj; device = device), kwargs...)
end
function QuantumStateSmoothPulseProblem(sys, ψ_init, ψ_goal, T, Δt; device::Int = 0, kwargs...)
    qtraj = KetTrajectory(sys, ψ_init, ψ_goal, T * Δt)
    traj = NamedTrajectory(qtraj, T)
    return SmoothPulseProblem(qtraj, T; integrator = B200BilinearIntegrator(qtraj, traj; device = device), kwargs...)
end
function UnitarySamplingProblem(systems, U_goal, T, Δt; device::Int = 0, kwargs...)
    qcp = UnitarySmoothPulseProblem(systems[1], U_goal, T, Δt; device = device)
    sqcp = SamplingProblem(qcp, systems; kwargs...)
    return replace_integrators!(sqcp, SamplingTrajectory(qcp.qtraj, systems); device = device)
end

end # module
