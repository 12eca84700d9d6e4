ssh test@127.0.0.1 "kill -9 3503"
ssh test@127.0.0.1 "kill -9 3504"
ssh test@127.0.0.1 "kill -9 3505"
kill -9 2946
