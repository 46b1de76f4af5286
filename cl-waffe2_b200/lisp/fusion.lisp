;;;; fusion.lisp — IR rewrite hook + CUDA-graph replay of a compiled iseq + data-parallel helpers.
;;;;
;;;; `on-finalizing-compiling` is called once per `build` per device with the forward / backward
;;;; InstructionSeq (`source/vm/nodes/defnode.lisp:26-57`, `vm/compile.lisp:237-238`; working
;;;; precedent: `source/backends/aten/tensor.lisp:105-141`).  Two rewrites are installed:
;;;;   1. MoveTensorNode immediately followed by an in-place {Add,Sub,Mul,Div}Node on the copy
;;;;      (`!add` = `!copy` + `A+=B`, `base-impl/arithmetic.lisp:150-152`) -> one wf2cu_ew_binary3 launch.
;;;;   2. Scalar->MatNode feeding a (:broadcast ..) operand of an arithmetic node
;;;;      (`fundamental.lisp:525-533`, `arithmetic.lisp:244-250`) -> wf2cu_ew_scalar with the host number
;;;;      as a kernel argument (no 4-byte H2D, SURVEY §7 hard part e).
;;;; Everything else is left as generated; each remaining instruction is already one kernel.
(in-package :cl-waffe2/backends.cuda)

(defun arith-op-code (node)
  (typecase node
    (AddNode +add+) (SubNode +sub+) (MulNode +mul+) (DivNode +div+) (t nil)))

(defun cuda-inst-p (inst)
  (every #'(lambda (x) (typep x 'CUDATensor)) (wf/vm:wfop-args inst)))

(defun fuse-move+arith (iseq)
  "copy <- MoveTensor(copy, x) ; copy <- Arith(copy, y)   ==>   copy <- binary3(op, copy, x, y)"
  (loop with out = nil
	while iseq
	do (let* ((i0 (pop iseq))
		  (i1 (car iseq)))
	     (if (and i1
		      (typep (wf/vm:wfop-node i0) 'MoveTensorNode)
		      (arith-op-code (wf/vm:wfop-node i1))
		      (cuda-inst-p i0) (cuda-inst-p i1)
		      (not (movetensor-ignore-me (wf/vm:wfop-node i0)))
		      (null (wf/vm:wfop-sv4bw i1))
		      ;; the arith node works in place on the Move's destination
		      (eq (first (wf/vm:wfop-out-to i0)) (first (wf/vm:wfop-args i1))))
		 (let ((code (arith-op-code (wf/vm:wfop-node i1))))
		   (pop iseq)
		   ;; keep i1's bookkeeping (out-to, node, self), replace its closure and arguments
		   (setf (wf/vm:wfop-args i1) (list (first (wf/vm:wfop-args i0))
						    (second (wf/vm:wfop-args i0))
						    (second (wf/vm:wfop-args i1)))
			 (wf/vm:wfop-op i1)
			 #'(lambda (dst x y)
			     (with-lowered-views ((dv dst) (xv x) (yv y))
			       (ew-binary3 code dv xv yv))
			     dst))
		   (push i1 out))
		 (push i0 out)))
	finally (return (nreverse out))))

(defmethod on-finalizing-compiling ((device-name (eql 'CUDATensor)) iseq-fw iseq-bw)
  (if *cuda-available-p*
      (values (fuse-move+arith iseq-fw) (and iseq-bw (fuse-move+arith iseq-bw)))
      (values iseq-fw iseq-bw)))

;; ---- CUDA-graph replay (SURVEY §8f rank 1): after `set-adjustable-symbols` the iseq is a fixed
;; launch sequence; capture `(forward model)` once and replay it.
(defmacro with-cuda-graph ((exec-var) capture-form &body body)
  `(let ((,exec-var (with-foreign-object (out :pointer)
		      (graph-begin)
		      ,capture-form
		      (graph-end out)
		      (mem-ref out :pointer))))
     (unwind-protect (progn ,@body)
       (graph-destroy ,exec-var))))

;; ---- data-parallel MLP step (SURVEY §8e): one process per GPU; gradients live in one arena
;; (allocate parameters' grads as views of a single CUDATensor) so the bucket is one range.
(defun cuda-comm-init (rank world id-bytes)
  "ID-BYTES: the 128-byte NCCL unique id made by rank 0 (wf2cu_comm_unique_id) and shared by file/env."
  (with-foreign-object (id :uint8 128)
    (dotimes (i 128) (setf (mem-aref id :uint8 i) (aref id-bytes i)))
    (comm-init rank world id)))

(defun cuda-allreduce-gradients (grad-arena n-floats)
  "Sum the gradient arena over ranks on the library stream; the 1/world factor is applied by the
optimizer kernel (`grad_scale` of wf2cu_adam_step)."
  (allreduce-sum-f32 (device-vec-ptr (tensor-vec grad-arena)) (tensor-initial-offset grad-arena) n-floats))
