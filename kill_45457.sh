ssh test@127.0.0.1 "kill -9 3674"
ssh test@127.0.0.1 "kill -9 3672"
ssh test@127.0.0.1 "kill -9 3673"
kill -9 2946
