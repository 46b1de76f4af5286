;;;; tensor.lisp — the CUDATensor device class: allocator, storage handle, vref, finalizer,
;;;; host<->device facet conversion.  Contract: `source/vm/generic-tensor/tensor.lisp:146-210`;
;;;; model: `source/backends/lisp/tensor.lisp:4-28`.
(in-package :cl-waffe2/backends.cuda)

(defstruct (device-vec (:constructor %make-device-vec (ptr bytes)))
  "The `vec` of a CUDATensor: a first-class handle because the VM shares storage with
(setf (tensor-vec a) (tensor-vec b)) (`source/vm/optimize-ir.lisp:4-18`, `allocation.lisp:523-525`).
The C side refcounts (wf2cu_retain/release); a GC finalizer releases handles that become garbage."
  (ptr (null-pointer) :type foreign-pointer)
  (bytes 0 :type fixnum))

(defun make-device-vec (n-elements &key (element-bytes 4))
  "wf2cu_malloc zero-fills, like LispTensor's `:initial-element 0` (`backends/lisp/tensor.lisp:20-28`);
`!sum` relies on it (`out * 0` must not see NaN, SURVEY §7 hard part b)."
  (with-foreign-object (out :pointer)
    (malloc (* n-elements element-bytes) out)
    (let* ((ptr (mem-ref out :pointer))
	   (vec (%make-device-vec ptr (* n-elements element-bytes))))
      (trivial-garbage:finalize vec #'(lambda () (%release ptr)))
      vec)))

(defclass CUDATensor (AbstractTensor) nil
  (:documentation "
## [AbstractTensor] CUDATensor
fp32 tensors resident in B200 HBM; every node runs as one kernel of libwaffe2cuda.so over the whole
strided view.  Enable with `(with-devices (CUDATensor LispTensor) ...)` or `(with-cuda ...)`."))

(defmethod current-backend-state ((backend-name (eql 'CUDATensor)))
  (if *cuda-available-p*
      (with-foreign-pointer-as-string (buf 128)
	(device-name buf 128))
      "not available (libwaffe2cuda.so / B200 missing)"))

(defmethod initialize-instance :before ((tensor CUDATensor) &rest initargs &key &allow-other-keys)
  (let* ((shape (getf initargs :shape))
	 (vec   (getf initargs :vec))
	 (facet (getf initargs :facet))
	 (initial-element (getf initargs :initial-element)))
    (when (eql facet :exist)
      (cond
	(vec (setf (tensor-vec tensor) vec))          ; adopt, do not allocate
	(t
	 (unless (eql (getf initargs :dtype) :float)
	   (error "CUDATensor supports :float only (ArgMax outputs are produced by the backend itself)"))
	 (let ((dv (make-device-vec (apply #'* shape))))
	   (setf (tensor-vec tensor) dv)
	   (when (and initial-element (/= initial-element 0))
	     (with-lowered-views ((v tensor)) ; lowering.lisp
	       (fill v (coerce initial-element 'single-float))))))))))

;; flat element access, no view (tensor.lisp:719-746).  Each call is a 4-byte transfer + sync:
;; used by Mat->Scalar / Scalar->Mat (`base-impl/fundamental.lisp:504-533`) and printing only.
(defmethod vref ((tensor CUDATensor) index)
  (with-foreign-object (out :float)
    (get-f32 (device-vec-ptr (tensor-vec tensor)) index out)
    (mem-ref out :float)))

(defmethod (setf vref) (new-value (tensor CUDATensor) index)
  (set-f32 (device-vec-ptr (tensor-vec tensor)) index (coerce new-value 'single-float))
  new-value)

(defmethod tensor-finalizer ((tensor CUDATensor))
  "Called when leaving (with-memory-pool ...) (`vm/allocation.lisp:108,170`)."
  (let ((vec (cl-waffe2/vm.generic-tensor::vec tensor)))
    (if (device-vec-p vec)
	#'(lambda ()
	    (trivial-garbage:cancel-finalization vec)
	    (%release (device-vec-ptr vec)))
	#'(lambda ()))))

;; ---- host <-> device (`source/array-converter.lisp:15-57,154-218`).  Bulk copies; the default
;; method would go element by element through vref.
(defmethod convert-tensor-facet ((from cl-waffe2/backends.lisp:LispTensor) (to (eql 'CUDATensor)))
  (let* ((src (if (cl-waffe2/vm.generic-tensor::permuted-p from) (->contiguous from) from))
	 (n   (apply #'* (cl-waffe2/vm.generic-tensor::original-shape src)))
	 (out (make-tensor (cl-waffe2/vm.generic-tensor::original-shape src) :dtype :float :device 'CUDATensor
			   :requires-grad (slot-value from 'cl-waffe2/vm.generic-tensor::requires-grad))))
    (with-pointer-to-vector-data (p (tensor-vec src))
      (h2d (device-vec-ptr (tensor-vec out)) 0 p (* 4 n)))
    out))

(defmethod convert-tensor-facet ((from CUDATensor) (to (eql 'cl-waffe2/backends.lisp:LispTensor)))
  (let* ((n   (floor (device-vec-bytes (tensor-vec from)) 4))
	 (out (make-tensor (cl-waffe2/vm.generic-tensor::original-shape from) :dtype :float
							      :device 'cl-waffe2/backends.lisp:LispTensor)))
    (with-pointer-to-vector-data (p (tensor-vec out))
      (d2h (device-vec-ptr (tensor-vec from)) 0 p (* 4 n)))   ; syncs
    out))

(defmethod convert-tensor-facet ((from CUDATensor) (to (eql 'simple-array)))
  (tensor-vec (convert-tensor-facet from 'cl-waffe2/backends.lisp:LispTensor)))

(defmacro with-cuda (&body body)
  "The `with-cuda` that `source/package.lisp:53` leaves commented out and `with-config :device :cuda`
expects (`source/utils.lisp:16,39-44`)."
  `(with-devices (CUDATensor cl-waffe2/backends.lisp:LispTensor)
     ,@body))

(defun cuda-sync () (sync))
